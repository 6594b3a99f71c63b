// synth_scene.cu — host side of the synthetic producer: scene and camera path from integer
// hashes, glm-compatible matrices (column-major, RH, depth 0..1; reference
// src/scene/camera.hpp:4-7,:27-41, src/main.cpp:93-108,:294,:329-336).
#include <cmath>
#include <cstring>
#include "common.cuh"
#include "../../include/vkr_synth.h"

namespace {

struct Pcg {  // PCG32 (XSH RR), O'Neill 2014
  uint64_t state, inc;
  explicit Pcg(uint64_t seed, uint64_t seq = 54u) : state(0), inc((seq << 1) | 1u) { next(); state += seed; next(); }
  uint32_t next() {
    uint64_t old = state;
    state = old * 6364136223846793005ULL + inc;
    uint32_t xs = (uint32_t)(((old >> 18u) ^ old) >> 27u), rot = (uint32_t)(old >> 59u);
    return (xs >> rot) | (xs << ((32u - rot) & 31u));
  }
  float unit() { return (float)(next() >> 8) * (1.0f / 16777216.0f); }  // [0,1), exact in fp32
  float range(float a, float b) { return a + (b - a) * unit(); }
};

uint32_t rgba(uint32_t r, uint32_t g, uint32_t b, uint32_t a = 255u) { return r | (g << 8) | (b << 16) | (a << 24); }

// sRGB codes whose decoded value is the roughness we want in material.g: ~0.05, 0.3, 0.6, 0.9
const uint32_t kRoughCode[4] = {63u, 149u, 203u, 243u};

void mat_mul(const double* A, const double* B, double* C) {  // column-major C = A*B
  for (int c = 0; c < 4; c++)
    for (int r = 0; r < 4; r++) {
      double s = 0;
      for (int k = 0; k < 4; k++) s += A[k * 4 + r] * B[c * 4 + k];
      C[c * 4 + r] = s;
    }
}

void look_at(const double* eye, const double* front, const double* up, double* V) {  // glm::lookAtRH
  double f[3] = {front[0], front[1], front[2]};
  double s[3] = {f[1] * up[2] - f[2] * up[1], f[2] * up[0] - f[0] * up[2], f[0] * up[1] - f[1] * up[0]};
  double sl = std::sqrt(s[0] * s[0] + s[1] * s[1] + s[2] * s[2]);
  for (int i = 0; i < 3; i++) s[i] /= sl;
  double u[3] = {s[1] * f[2] - s[2] * f[1], s[2] * f[0] - s[0] * f[2], s[0] * f[1] - s[1] * f[0]};
  double M[16] = {s[0], u[0], -f[0], 0, s[1], u[1], -f[1], 0, s[2], u[2], -f[2], 0,
                  -(s[0] * eye[0] + s[1] * eye[1] + s[2] * eye[2]), -(u[0] * eye[0] + u[1] * eye[1] + u[2] * eye[2]),
                  (f[0] * eye[0] + f[1] * eye[1] + f[2] * eye[2]), 1};
  memcpy(V, M, sizeof(M));
}

void camera_pose(uint32_t view_seed, uint32_t frame, double* V) {
  Pcg rng(0xC0FFEEull + view_seed, 7u);
  double ex = rng.range(-3.0f, 3.0f), ey = rng.range(1.2f, 2.2f), ez = rng.range(-7.0f, -4.0f);
  double yaw0 = 90.0 + rng.range(-20.0f, 20.0f), pitch = rng.range(2.0f, 12.0f);  // world_up = -y => positive pitch looks down
  if (view_seed == 0) { ex = 0.0; ey = 1.6; ez = -6.0; yaw0 = 90.0; pitch = 6.0; }
  const double deg = 3.14159265358979323846 / 180.0;
  double yaw = yaw0 + 0.1 * frame;
  // scene::Camera::update_camera_vectors (reference src/scene/camera.hpp:31-41) with world_up = (0,-1,0);
  // pitch sign flipped so that positive = toward the ground in a y-up world.
  double fr[3] = {std::cos(yaw * deg) * std::cos(pitch * deg), -std::sin(pitch * deg), std::sin(yaw * deg) * std::cos(pitch * deg)};
  double wu[3] = {0, -1, 0};
  double r[3] = {fr[1] * wu[2] - fr[2] * wu[1], fr[2] * wu[0] - fr[0] * wu[2], fr[0] * wu[1] - fr[1] * wu[0]};
  double rl = std::sqrt(r[0] * r[0] + r[1] * r[1] + r[2] * r[2]);
  for (int i = 0; i < 3; i++) r[i] /= rl;
  double up[3] = {r[1] * fr[2] - r[2] * fr[1], r[2] * fr[0] - r[0] * fr[2], r[0] * fr[1] - r[1] * fr[0]};
  // translate 0.02 units/frame along the initial view direction projected on the ground
  double f0[2] = {std::cos(yaw0 * deg), std::sin(yaw0 * deg)};
  double eye[3] = {ex + 0.02 * frame * f0[0], ey, ez + 0.02 * frame * f0[1]};
  look_at(eye, fr, up, V);
}

void perspective(double fovy, double aspect, double zn, double zf, double* P) {  // glm::perspectiveRH_ZO
  double t = std::tan(fovy / 2.0);
  memset(P, 0, 16 * sizeof(double));
  P[0] = 1.0 / (aspect * t); P[5] = 1.0 / t; P[10] = zf / (zn - zf); P[11] = -1.0; P[14] = -(zf * zn) / (zf - zn);
}

}  // namespace

extern "C" void vkr_synth_scene(uint32_t seed, uint32_t object_count, VkrSynthScene* out) {
  memset(out, 0, sizeof(*out));
  if (object_count > VKR_SYNTH_MAX_OBJECTS) object_count = VKR_SYNTH_MAX_OBJECTS;
  Pcg rng(0x5EED0000ull + seed, 11u);
  out->object_count = object_count;
  out->wall_z = 9.0f; out->wall_height = 5.0f; out->wall_half_width = 14.0f;
  out->ground_albedo[0] = rgba(200, 196, 188); out->ground_albedo[1] = rgba(96, 104, 118);
  out->ground_material = rgba(0, kRoughCode[1], 0);
  out->wall_albedo = rgba(178, 121, 88); out->wall_material = rgba(0, kRoughCode[2], 0);
  out->sky_albedo = rgba(120, 160, 220);
  for (uint32_t i = 0; i < object_count; i++) {
    VkrSynthObject& o = out->objects[i];
    o.kind = rng.next() & 1u;
    float sx = rng.range(0.25f, 0.9f), sy = rng.range(0.25f, 1.1f), sz = rng.range(0.25f, 0.9f);
    o.half_extent[0] = sx; o.half_extent[1] = o.kind ? sy : sx; o.half_extent[2] = o.kind ? sz : sx;
    o.center[0] = rng.range(-7.5f, 7.5f);
    o.center[2] = rng.range(-2.5f, 8.0f);
    o.center[1] = o.half_extent[1] + ((rng.next() & 7u) == 0u ? rng.range(0.2f, 1.5f) : 0.0f);  // some float above the ground
    o.albedo_srgb8 = rgba(64 + (rng.next() % 192u), 64 + (rng.next() % 192u), 64 + (rng.next() % 192u));
    o.material_srgb8 = rgba(0, kRoughCode[rng.next() & 3u], (rng.next() & 1u) ? 255u : 0u);
  }
}

extern "C" void vkr_synth_camera(uint32_t view_seed, uint32_t frame, uint32_t width, uint32_t height, int use_jitter, VkrDrawTAAParams* out) {
  const double fovy = 60.0 * 3.14159265358979323846 / 180.0, zn = 0.05, zf = 80.0;
  const double aspect = (double)((float)width / (float)height);
  double V[16], Vp[16], P[16], M[16], Mp[16];
  camera_pose(view_seed, frame, V);
  camera_pose(view_seed, frame ? frame - 1 : 0, Vp);
  perspective(fovy, aspect, zn, zf, P);
  mat_mul(P, V, M);
  mat_mul(P, Vp, Mp);
  for (int i = 0; i < 16; i++) {
    out->mvp[i] = (float)M[i]; out->prev_mvp[i] = (float)Mp[i];
    out->camera[i] = (float)V[i]; out->prev_camera[i] = (float)Vp[i];
  }
  // reference src/main.cpp:93-108 next_taa_offset: (2*o - 1) / resolution, 4-sample cycle
  static const float offs[4][2] = {{0.25f, 0.25f}, {0.75f, 0.75f}, {0.75f, 0.25f}, {0.25f, 0.75f}};
  out->jitter[0] = use_jitter ? (2.0f * offs[frame & 3][0] - 1.0f) * (float)(1.0 / width) : 0.f;
  out->jitter[1] = use_jitter ? (2.0f * offs[frame & 3][1] - 1.0f) * (float)(1.0 / height) : 0.f;
  out->jitter[2] = out->jitter[3] = 0.f;
  out->fovy_aspect_znear_zfar[0] = (float)fovy;  // glm::radians(60.f)
  out->fovy_aspect_znear_zfar[1] = (float)width / (float)height;
  out->fovy_aspect_znear_zfar[2] = 0.05f; out->fovy_aspect_znear_zfar[3] = 80.f;
}

// General 4x4 inverse (cofactors, evaluated in double) — stands in for glm::inverse.
extern "C" void vkr_mat4_inverse(const float* mf, float* out) {
  double m[16], inv[16];
  for (int i = 0; i < 16; i++) m[i] = mf[i];
  inv[0] = m[5] * m[10] * m[15] - m[5] * m[11] * m[14] - m[9] * m[6] * m[15] + m[9] * m[7] * m[14] + m[13] * m[6] * m[11] - m[13] * m[7] * m[10];
  inv[4] = -m[4] * m[10] * m[15] + m[4] * m[11] * m[14] + m[8] * m[6] * m[15] - m[8] * m[7] * m[14] - m[12] * m[6] * m[11] + m[12] * m[7] * m[10];
  inv[8] = m[4] * m[9] * m[15] - m[4] * m[11] * m[13] - m[8] * m[5] * m[15] + m[8] * m[7] * m[13] + m[12] * m[5] * m[11] - m[12] * m[7] * m[9];
  inv[12] = -m[4] * m[9] * m[14] + m[4] * m[10] * m[13] + m[8] * m[5] * m[14] - m[8] * m[6] * m[13] - m[12] * m[5] * m[10] + m[12] * m[6] * m[9];
  inv[1] = -m[1] * m[10] * m[15] + m[1] * m[11] * m[14] + m[9] * m[2] * m[15] - m[9] * m[3] * m[14] - m[13] * m[2] * m[11] + m[13] * m[3] * m[10];
  inv[5] = m[0] * m[10] * m[15] - m[0] * m[11] * m[14] - m[8] * m[2] * m[15] + m[8] * m[3] * m[14] + m[12] * m[2] * m[11] - m[12] * m[3] * m[10];
  inv[9] = -m[0] * m[9] * m[15] + m[0] * m[11] * m[13] + m[8] * m[1] * m[15] - m[8] * m[3] * m[13] - m[12] * m[1] * m[11] + m[12] * m[3] * m[9];
  inv[13] = m[0] * m[9] * m[14] - m[0] * m[10] * m[13] - m[8] * m[1] * m[14] + m[8] * m[2] * m[13] + m[12] * m[1] * m[10] - m[12] * m[2] * m[9];
  inv[2] = m[1] * m[6] * m[15] - m[1] * m[7] * m[14] - m[5] * m[2] * m[15] + m[5] * m[3] * m[14] + m[13] * m[2] * m[7] - m[13] * m[3] * m[6];
  inv[6] = -m[0] * m[6] * m[15] + m[0] * m[7] * m[14] + m[4] * m[2] * m[15] - m[4] * m[3] * m[14] - m[12] * m[2] * m[7] + m[12] * m[3] * m[6];
  inv[10] = m[0] * m[5] * m[15] - m[0] * m[7] * m[13] - m[4] * m[1] * m[15] + m[4] * m[3] * m[13] + m[12] * m[1] * m[7] - m[12] * m[3] * m[5];
  inv[14] = -m[0] * m[5] * m[14] + m[0] * m[6] * m[13] + m[4] * m[1] * m[14] - m[4] * m[2] * m[13] - m[12] * m[1] * m[6] + m[12] * m[2] * m[5];
  inv[3] = -m[1] * m[6] * m[11] + m[1] * m[7] * m[10] + m[5] * m[2] * m[11] - m[5] * m[3] * m[10] - m[9] * m[2] * m[7] + m[9] * m[3] * m[6];
  inv[7] = m[0] * m[6] * m[11] - m[0] * m[7] * m[10] - m[4] * m[2] * m[11] + m[4] * m[3] * m[10] + m[8] * m[2] * m[7] - m[8] * m[3] * m[6];
  inv[11] = -m[0] * m[5] * m[11] + m[0] * m[7] * m[9] + m[4] * m[1] * m[11] - m[4] * m[3] * m[9] - m[8] * m[1] * m[7] + m[8] * m[3] * m[5];
  inv[15] = m[0] * m[5] * m[10] - m[0] * m[6] * m[9] - m[4] * m[1] * m[10] + m[4] * m[2] * m[9] + m[8] * m[1] * m[6] - m[8] * m[2] * m[5];
  double det = m[0] * inv[0] + m[1] * inv[4] + m[2] * inv[8] + m[3] * inv[12];
  double id = det != 0.0 ? 1.0 / det : 0.0;
  for (int i = 0; i < 16; i++) out[i] = (float)(inv[i] * id);
}

extern "C" void vkr_mat4_transpose(const float* m, float* out) {
  float t[16];
  for (int c = 0; c < 4; c++)
    for (int r = 0; r < 4; r++) t[r * 4 + c] = m[c * 4 + r];
  memcpy(out, t, sizeof(t));
}
