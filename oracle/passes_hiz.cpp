// ORACLE — TEST INFRASTRUCTURE ONLY (see oracle.h header).  PARITY UNPINNED.
//
// Hi-Z min-depth pyramid + half-res G-buffer, Halton table, LUT pre-integration.
#include <omp.h>
#include <cmath>
#include <mutex>
#include "oracle.h"
#include "shader_include.hpp"

namespace orc {

// sRGB transfer functions, built once in double precision (the product builds the same
// table on its host side; Vulkan 1.2 spec "sRGB EOTF").
static float g_srgb_table[256];
static std::once_flag g_srgb_once;
const float* srgb_decode_table() {
  std::call_once(g_srgb_once, [] {
    for (int i = 0; i < 256; i++) {
      double c = i / 255.0;
      double l = (c <= 0.04045) ? c / 12.92 : std::pow((c + 0.055) / 1.055, 2.4);
      g_srgb_table[i] = (float)l;
    }
  });
  return g_srgb_table;
}
float srgb_encode(float l) {
  if (!(l > 0.0f)) return 0.0f;
  if (l >= 1.0f) return 1.0f;
  if (l < 0.0031308f) return 12.92f * l;
  return 1.055f * std::pow(l, 1.0f / 2.4f) - 0.055f;
}

static inline void row_range(int h, int& y0, int& y1) { if (y0 < 0) y0 = 0; if (y1 <= 0 || y1 > h) y1 = h; }
static inline uint32_t load_u32(const Image& im, int x, int y, int lod) { uint32_t v; std::memcpy(&v, texel_ptr(im, x, y, lod), 4); return v; }
static inline void store_u32(const Image& im, int x, int y, int lod, uint32_t v) { std::memcpy((void*)texel_ptr(im, x, y, lod), &v, 4); }

// D24 texel fetch as integer; out-of-range = 0 (App. A.3).  The comparison of the shader's
// float(k)/16777215 values is order-isomorphic to comparing k (App. A.5).
static inline uint32_t fetch_d24(const Image& im, int x, int y, int lod) {
  int w = mip_dim(im.width, lod), h = mip_dim(im.height, lod);
  if (x < 0 || y < 0 || x >= w || y >= h) return 0u;
  return load_u32(im, x, y, lod) & 0xFFFFFFu;
}

}  // namespace orc
using namespace orc;

extern "C" {

void orc_set_threads(int n) { if (n > 0) omp_set_num_threads(n); }
int orc_get_max_threads(void) { return omp_get_max_threads(); }

// reference src/advanced_ssr.cpp:8-20
static float halton_elem(uint32_t index, uint32_t base) {
  float f = 1;
  float r = 0;
  uint32_t current = index;
  do {
    f = f / base;
    r = r + f * (current % base);
    current = (uint32_t)std::floor(current / float(base));
  } while (current > 0);
  return r;
}

// reference src/advanced_ssr.cpp:22-34
void orc_halton23_seq(uint32_t count, float* out) {
  for (uint32_t iter = 0; iter < count; iter++) {
    out[4 * iter + 0] = halton_elem(iter + 1, 2);
    out[4 * iter + 1] = halton_elem(iter + 1, 3);
    out[4 * iter + 2] = 0; out[4 * iter + 3] = 0;
  }
}

// reference shaders/advanced_ssr/downsample_gbuffer.frag:12-37; host src/downsample_pass.cpp:25-92
int orc_downsample_gbuffer(const OrcImage* depth_, const OrcImage* normal_, const OrcImage* velocity_,
                           const OrcImage* out_normal_, const OrcImage* out_velocity_, int y0, int y1) {
  const Image& depth = *(const Image*)depth_; const Image& normal = *(const Image*)normal_;
  const Image& velocity = *(const Image*)velocity_; const Image& on = *(const Image*)out_normal_; const Image& ov = *(const Image*)out_velocity_;
  if (depth.mip_count < 2) return -1;                                     // downsample_pass.cpp:37-39
  int w = max(1, (int)depth.width / 2), h = max(1, (int)depth.height / 2);
  if ((int)on.width != w || (int)on.height != h || (int)ov.width != w || (int)ov.height != h) return -2;  // :48-50
  row_range(h, y0, y1);
#pragma omp parallel for schedule(dynamic, 8)
  for (int y = y0; y < y1; y++)
    for (int x = 0; x < w; x++) {
      int px = 2 * x, py = 2 * y;
      uint32_t d0 = fetch_d24(depth, px, py, 0), d1 = fetch_d24(depth, px + 1, py, 0);
      uint32_t d2 = fetch_d24(depth, px, py + 1, 0), d3 = fetch_d24(depth, px + 1, py + 1, 0);
      int ox = 0, oy = 0;
      uint32_t min_depth = std::min(std::min(d0, d1), std::min(d2, d3));
      if (min_depth == d1) { ox = 1; oy = 0; }
      else if (min_depth == d2) { ox = 0; oy = 1; }
      else if (min_depth == d3) { ox = 1; oy = 1; }
      // texelFetch -> attachment store of an already representable value is the identity
      // on the raw bits; out-of-range source = 0.
      int sx = px + ox, sy = py + oy;
      uint32_t nbits = (sx < (int)normal.width && sy < (int)normal.height) ? load_u32(normal, sx, sy, 0) : 0u;
      uint32_t vbits = (sx < (int)velocity.width && sy < (int)velocity.height) ? load_u32(velocity, sx, sy, 0) : 0u;
      store_u32(on, x, y, 0, nbits);
      store_u32(ov, x, y, 0, vbits);
      store_u32(depth, x, y, 1, min_depth);   // gl_FragDepth, stencil byte = 0
    }
  return 0;
}

// reference shaders/advanced_ssr/depth_mips.frag:7-15; host src/downsample_pass.cpp:94-131
int orc_depth_mips(const OrcImage* depth_, uint32_t src_mip) {
  const Image& depth = *(const Image*)depth_;
  for (uint32_t i = src_mip + 1; i < depth.mip_count; i++) {
    int w = mip_dim(depth.width, i), h = mip_dim(depth.height, i);
#pragma omp parallel for schedule(dynamic, 8)
    for (int y = 0; y < h; y++)
      for (int x = 0; x < w; x++) {
        int px = 2 * x, py = 2 * y;
        uint32_t d0 = fetch_d24(depth, px, py, i - 1), d1 = fetch_d24(depth, px + 1, py, i - 1);
        uint32_t d2 = fetch_d24(depth, px, py + 1, i - 1), d3 = fetch_d24(depth, px + 1, py + 1, i - 1);
        store_u32(depth, x, y, i, std::min(std::min(d0, d1), std::min(d2, d3)));
      }
  }
  return 0;
}

// reference shaders/advanced_ssr/preintegrate.comp:45-65,:77-84
static float G2(float t, float a, float b) {
  const float p = b - a;
  const float q = b + a;
  const float L = p * t + q;
  const float nom = (1 - t) * L;
  const float denom = (1 + t * t - 0.5f * L * L);
  return (L > 0) ? nom / (denom * denom) : 0.0f;
}

int orc_pdf_preintegrate(const OrcImage* out_, int y0, int y1) {
  const Image& out = *(const Image*)out_;
  const int STEP_COUNT = 2000;
  int rw = out.width, rh = out.height;
  row_range(rh, y0, y1);
#pragma omp parallel for schedule(dynamic, 4)
  for (int y = y0; y < y1; y++)
    for (int x = 0; x < rw; x++) {
      const float a = 2 * float(x + 0.5f) / rw - 1;
      const float b = float(y + 0.5f) / rh;
      float sum = 0.0f;
      float dt = 2.0f / STEP_COUNT;
      for (int i = 0; i < STEP_COUNT; i++) {
        float t = -1 + dt * (i + 0.5f);
        sum += G2(t, a, b);
      }
      float g = 2.0f / STEP_COUNT * sum;
      imageStore(out, 0, ivec2(x, y), vec4(g, 0, 0, 0));
    }
  return 0;
}

// reference shaders/advanced_ssr/preintegrate_ssr.comp:12-45
int orc_brdf_preintegrate(const float* samples, const OrcImage* out_, int y0, int y1) {
  const Image& out = *(const Image*)out_;
  const int NUM_SAMPLES = 128;
  int tw = out.width, th = out.height;
  row_range(th, y0, y1);
#pragma omp parallel for schedule(dynamic, 4)
  for (int y = y0; y < y1; y++)
    for (int x = 0; x < tw; x++) {
      const float roughness = (x + 0.5f) / float(tw);
      const float NdotV = (y + 0.5f) / float(th);
      const float roughness2 = roughness * roughness;
      const vec3 V = vec3(sqrt(1 - NdotV * NdotV), 0, NdotV);
      float A = 0.0f;
      float B = 0.0f;
      for (int i = 0; i < NUM_SAMPLES; i++) {
        vec3 H = sampleGGXVNDF(V, roughness2, roughness2, samples[4 * i], samples[4 * i + 1]);
        vec3 L = normalize(reflect(-V, H));
        float NdotL = L.z;
        float alpha = std::pow(1 - dot(V, H), 5.0f);
        float G1 = brdfG1(roughness2, NdotV);
        float G2v = brdfG2(NdotV, NdotL, roughness2);
        A += G2v / G1 * (1 - alpha);
        B += G2v / G1 * alpha;
      }
      A *= 1.0f / NUM_SAMPLES;
      B *= 1.0f / NUM_SAMPLES;
      imageStore(out, 0, ivec2(x, y), vec4(A, B, 0, 0));
    }
  return 0;
}

void orc_probe_encode_normal(const float* v, float* o) { vec2 r = encode_normal(vec3(v[0], v[1], v[2])); o[0] = r.x; o[1] = r.y; }
void orc_probe_decode_normal(const float* uv, float* o) { vec3 r = decode_normal(vec2(uv[0], uv[1])); o[0] = r.x; o[1] = r.y; o[2] = r.z; }
void orc_probe_reconstruct(const float* uv, float d, const float* cam, float* o) {
  vec3 r = reconstruct_view_vec(vec2(uv[0], uv[1]), d, Cam(cam[0], cam[1], cam[2], cam[3])); o[0] = r.x; o[1] = r.y; o[2] = r.z; }
void orc_probe_project(const float* v, const float* cam, float* o) {
  vec3 r = project_view_vec(vec3(v[0], v[1], v[2]), Cam(cam[0], cam[1], cam[2], cam[3])); o[0] = r.x; o[1] = r.y; o[2] = r.z; }
float orc_probe_half_roundtrip(float v) { return half_to_float(float_to_half(v)); }
float orc_probe_srgb_decode(uint32_t c) { return srgb_decode_table()[c & 255]; }
float orc_probe_srgb_encode(float l) { return srgb_encode(l); }

}  // extern "C"
