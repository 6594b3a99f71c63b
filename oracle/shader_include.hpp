// ORACLE — TEST INFRASTRUCTURE ONLY (see glsl.hpp header).
//
// Restatement of the shared shader includes, function by function, same names and
// statement order so they can be diffed by eye against the GLSL:
//   reference src/shaders/include/gbuffer_encode.glsl:5-84
//   reference src/shaders/include/brdf.glsl:4-13,:31-56,:107-155
//   reference src/shaders/include/screen_trace.glsl:4-49
#pragma once
#include "image.hpp"

namespace orc {

// ---- gbuffer_encode.glsl ---------------------------------------------------
inline float sign_nz(float k) { return (k >= 0.0f) ? 1.0f : -1.0f; }                     // :5-7
inline vec2 sign_nz(vec2 v) { return vec2(sign_nz(v.x), sign_nz(v.y)); }                 // :9-11

inline vec2 encode_normal(vec3 v) {                                                      // :17-27
  float l1norm = abs(v.x) + abs(v.y) + abs(v.z);
  vec2 result = v.xy() * (1.0f / l1norm);
  if (v.z < 0.0f) {
    result = (1.0f - abs(vec2(result.y, result.x))) * sign_nz(result);
  }
  return 0.5f * result + vec2(0.5f, 0.5f);
}

inline vec3 decode_normal(vec2 uv) {                                                     // :29-37
  uv = 2.f * uv - vec2(1.f, 1.f);
  vec3 v = vec3(uv.x, uv.y, 1.0f - abs(uv.x) - abs(uv.y));
  if (v.z < 0.0f) {
    vec2 t = (1.0f - abs(vec2(v.y, v.x))) * sign_nz(v.xy());
    v.x = t.x; v.y = t.y;
  }
  return normalize(v);
}

inline vec3 sample_gbuffer_normal(const View& normal_tex, vec2 uv) {                     // :39-43
  vec2 t = texture(normal_tex, uv).xy();
  return decode_normal(t);
}

inline float linearize_depth2(float d, float n, float f) {                               // :53-56
  return n * f / (d * (f - n) - f);
}

// Camera constants bundle.  The GLSL recomputes tan(fovy/2) per call (:60,:76); it is a
// pure function of a uniform, so it is evaluated once here with the host's tanf — the
// product's host side does the same and passes the identical float to its kernels.
struct Cam {
  float fovy, aspect, znear, zfar, tg_alpha;
  Cam(float fy, float a, float n, float f) : fovy(fy), aspect(a), znear(n), zfar(f), tg_alpha(std::tan(fy / 2)) {}
};

inline vec3 reconstruct_view_vec(vec2 uv, float d, const Cam& c) {                       // :58-69
  float tg_alpha = c.tg_alpha;
  float z = linearize_depth2(d, c.znear, c.zfar);
  float xd = 2 * uv.x - 1;
  float yd = 2 * uv.y - 1;
  float x = -(xd) * (z * c.aspect * tg_alpha);
  float y = -(yd) * (z * tg_alpha);
  return vec3(x, y, z);
}

inline float encode_depth(float z, float n, float f) {                                   // :71-73
  return f / (f - n) + f * n / (z * (f - n));
}

inline vec3 project_view_vec(vec3 v, const Cam& c) {                                     // :75-84
  float tg_alpha = c.tg_alpha;
  float n = c.znear, f = c.zfar;
  float z = v.z;
  float depth = f / (f - n) + f * n / (z * (f - n));
  float pu = v.x / (-v.z * tg_alpha * c.aspect);
  float pv = v.y / (-z * tg_alpha);
  return vec3(0.5f * pu + 0.5f, 0.5f * pv + 0.5f, depth);
}

// ---- brdf.glsl -------------------------------------------------------------
inline vec3 fresnelSchlick(float cos_theta, vec3 F0) {                                   // :6-8
  return F0 + (1.0f - F0) * std::pow(clamp(1.0f - cos_theta, 0.0f, 1.0f), 5.0f);
}

inline vec3 F0_approximation(vec3 albedo, float metallic) {                              // :10-13
  vec3 F0 = vec3(0.04f);
  return mix(F0, albedo, metallic);
}

inline float DistributionGGX(vec3 N, vec3 H, float alpha) {                              // :31-38
  float NoH = dot(N, H);
  float alpha2 = alpha * alpha;
  float NoH2 = NoH * NoH;
  float den = NoH2 * alpha2 + (1 - NoH2);
  return (((NoH2 > 0) ? 1 : 0) * alpha2) / (PI * den * den);
}

inline float brdfG1(float alpha2, float NdotV) {                                         // :43-47
  float NdotV2 = NdotV * NdotV;
  float tgv2 = (1 - NdotV2) / NdotV2;
  return 2.0f / (1 + sqrt(1 + alpha2 * tgv2));
}

inline float brdfG2(float NdotV, float NdotL, float alpha2) {                            // :49-56
  float NdotV2 = NdotV * NdotV;
  float NdotL2 = NdotL * NdotL;
  float L1 = sqrt(1 + alpha2 * (1 - NdotV2) / NdotV2);
  float L2 = sqrt(1 + alpha2 * (1 - NdotL2) / NdotL2);
  return 2.0f / (L1 + L2);
}

inline float sampleGGXdirPDF(const View& PDF_TEX, vec3 V, vec3 N, vec3 L, float alpha) { // :107-129
  vec3 Y = normalize(cross(V, N));
  vec3 X = normalize(cross(Y, V));
  alpha = clamp(alpha, 0, 0.9f);
  vec3 Lproj = normalize(L - V * dot(V, L));
  float cos_theta = dot(X, Lproj);
  const float cos_phin = dot(N, X);
  const float sin_phin = sqrt(1 - cos_phin * cos_phin);
  const float alpha2 = alpha * alpha;
  const float coef = sqrt(1 - alpha2);
  const float a = 0.5f * coef * cos_phin * cos_theta + 0.5f;
  const float b = coef * sin_phin;
  float pdf = alpha2 / (2 * PI * coef) * texture(PDF_TEX, vec2(a, b)).x;
  return pdf;
}

inline vec3 sampleGGXVNDF(vec3 Ve, float alpha_x, float alpha_y, float U1, float U2) {   // :135-155
  vec3 Vh = normalize(vec3(alpha_x * Ve.x, alpha_y * Ve.y, Ve.z));
  float lensq = Vh.x * Vh.x + Vh.y * Vh.y;
  vec3 T1 = lensq > 0 ? vec3(-Vh.y, Vh.x, 0) * inversesqrt(lensq) : vec3(1, 0, 0);
  vec3 T2 = cross(Vh, T1);
  float r = sqrt(U1);
  float phi = 2.0f * PI * U2;
  float t1 = r * std::cos(phi);
  float t2 = r * std::sin(phi);
  float s = 0.5f * (1.0f + Vh.z);
  t2 = (1.0f - s) * sqrt(1.0f - t1 * t1) + s * t2;
  vec3 Nh = t1 * T1 + t2 * T2 + sqrt(max(0.0f, 1.0f - t1 * t1 - t2 * t2)) * Vh;
  vec3 Ne = normalize(vec3(alpha_x * Nh.x, alpha_y * Nh.y, max(0.0f, Nh.z)));
  return Ne;
}

// ---- screen_trace.glsl -----------------------------------------------------
constexpr float MAX_T_FLOAT = 3.402823466e+38f;                                          // :4

inline void initial_advance_ray(vec3 origin, vec3 dir, vec3 inv_dir, vec2 mip_res, vec2 inv_mip_res,
                                vec2 floor_offset, vec2 uv_offset, vec3& pos, float& current_t) {  // :8-15
  vec2 cur_pos = mip_res * origin.xy();
  vec2 xy_plane = floor(cur_pos) + floor_offset;
  xy_plane = xy_plane * inv_mip_res + uv_offset;
  vec2 t = (xy_plane - origin.xy()) * inv_dir.xy();
  current_t = min(t.x, t.y);
  pos = origin + current_t * dir;
}

inline bool advance_ray(vec3 origin, vec3 direction, vec3 inv_direction, vec2 current_mip_position,
                        vec2 current_mip_resolution_inv, vec2 floor_offset, vec2 uv_offset, float surface_z,
                        vec3& position, float& current_t) {                              // :17-45
  vec2 xy_plane = floor(current_mip_position) + floor_offset;
  xy_plane = xy_plane * current_mip_resolution_inv + uv_offset;
  vec3 boundary_planes = vec3(xy_plane, surface_z);
  vec3 t = (boundary_planes - origin) * inv_direction;
  t.z = direction.z > 0 ? t.z : MAX_T_FLOAT;
  float t_min = min(min(t.x, t.y), t.z);
  bool above_surface = surface_z > position.z;
  bool skipped_tile = t_min != t.z && above_surface;
  current_t = above_surface ? t_min : current_t;
  position = origin + current_t * direction;
  return skipped_tile;
}

inline vec2 get_mip_resolution(vec2 screen_dimensions, int mip_level) {                  // :47-49
  return screen_dimensions * std::pow(0.5f, float(mip_level));
}

// screen_trace.glsl:51-100
inline vec3 hierarchical_raymarch(const View& depth_tex, vec3 origin, vec3 direction, int most_detailed_mip, uint32_t max_traversal_intersections, bool& valid_hit) {
  const vec3 inv_direction = vec3(direction.x != 0 ? 1.0f / direction.x : MAX_T_FLOAT, direction.y != 0 ? 1.0f / direction.y : MAX_T_FLOAT,
                                  direction.z != 0 ? 1.0f / direction.z : MAX_T_FLOAT);
  int current_mip = most_detailed_mip;
  vec2 screen_size = to_vec2(textureSize(depth_tex, 0));
  vec2 current_mip_resolution = get_mip_resolution(screen_size, current_mip);
  vec2 current_mip_resolution_inv = 1.0f / current_mip_resolution;
  vec2 uv_offset = 0.005f * std::exp2((float)most_detailed_mip) / screen_size;
  uv_offset.x = direction.x < 0 ? -uv_offset.x : uv_offset.x;
  uv_offset.y = direction.y < 0 ? -uv_offset.y : uv_offset.y;
  vec2 floor_offset = vec2(direction.x < 0 ? 0 : 1, direction.y < 0 ? 0 : 1);
  float current_t;
  vec3 position;
  initial_advance_ray(origin, direction, inv_direction, current_mip_resolution, current_mip_resolution_inv, floor_offset, uv_offset, position, current_t);
  uint32_t i = 0;
  while (i < max_traversal_intersections && current_mip >= most_detailed_mip) {
    vec2 current_mip_position = current_mip_resolution * position.xy();
    float surface_z = texelFetch(depth_tex, to_ivec2(current_mip_position), current_mip).x;
    bool skipped_tile = advance_ray(origin, direction, inv_direction, current_mip_position, current_mip_resolution_inv, floor_offset, uv_offset, surface_z, position, current_t);
    current_mip += skipped_tile ? 1 : -1;
    current_mip_resolution *= skipped_tile ? 0.5f : 2.0f;
    current_mip_resolution_inv *= skipped_tile ? 2.0f : 0.5f;
    ++i;
  }
  valid_hit = (i <= max_traversal_intersections);
  return position;
}

}  // namespace orc
