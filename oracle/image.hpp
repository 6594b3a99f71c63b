// ORACLE — TEST INFRASTRUCTURE ONLY (see glsl.hpp header).
//
// Image formats, texel codecs and the one sampler the hot path uses
// (gpu::DEFAULT_SAMPLER, reference src/gpu/samplers.hpp:36-55: LINEAR min/mag,
// CLAMP_TO_EDGE).  Pinned interpretation — SURVEY.md App. A.1/A.3/A.5:
//   * D24_UNORM_S8_UINT texel = u32, depth in the low 24 bits (readback layout,
//     reference src/main.cpp:139); shader value = float(k) / 16777215.0f.
//   * bilinear: x = u*W - 0.5, i0 = floor(x), w = x - i0 (fp32), indices clamped to
//     the edge, sRGB decoded before weighting, mix() in x then in y.
//   * texelFetch out of range (coordinate or lod) = all zero.
//   * float -> unorm store: clamp to [0,1] (NaN -> 0), floor(v*max + 0.5).
//   * float -> sfloat16 store: IEEE round-to-nearest-even.
#pragma once
#include <cstdint>
#include <cstring>
#include "glsl.hpp"

namespace orc {
using namespace glsl;

// Numeric values = VkFormat, so the descriptor reads like the reference's ImageInfo.
enum Format : uint32_t {
  FMT_R8_UNORM = 9,
  FMT_R8G8B8A8_UNORM = 37,
  FMT_R8G8B8A8_SRGB = 43,
  FMT_R16_UNORM = 70,
  FMT_R16_SFLOAT = 76,
  FMT_R16G16_UNORM = 77,
  FMT_R16G16_SFLOAT = 83,
  FMT_R16G16B16A16_UNORM = 91,
  FMT_R16G16B16A16_SFLOAT = 97,
  FMT_R32_SFLOAT = 100,
  FMT_R32G32B32A32_SFLOAT = 109,
  FMT_D24_UNORM_S8_UINT = 129,
};

// Same POD layout as VkrImage in include/vkr_passes.h (kept separate on purpose:
// the oracle includes no product header).
struct Image {
  void* base;
  uint32_t width, height;
  uint32_t format;
  uint32_t mip_count;
  uint64_t mip_offset[16];  // bytes from base
  uint32_t row_pitch[16];   // bytes
};

inline uint32_t texel_size(uint32_t f) {
  switch (f) {
    case FMT_R8_UNORM: return 1;
    case FMT_R16_UNORM: case FMT_R16_SFLOAT: return 2;
    case FMT_R16G16B16A16_UNORM: case FMT_R16G16B16A16_SFLOAT: return 8;
    case FMT_R32G32B32A32_SFLOAT: return 16;
    default: return 4;
  }
}

inline int mip_dim(uint32_t d, int lod) { uint32_t v = d >> lod; return v ? (int)v : 1; }

// ---- scalar codecs -------------------------------------------------------
inline float half_to_float(uint16_t h) {
  uint32_t s = (uint32_t)(h >> 15) << 31, e = (h >> 10) & 31, m = h & 1023, bits;
  if (e == 0) {
    if (m == 0) bits = s;
    else { int sh = 0; while (!(m & 1024)) { m <<= 1; sh++; } m &= 1023; bits = s | ((uint32_t)(113 - sh) << 23) | (m << 13); }
  } else if (e == 31) bits = s | 0x7F800000u | (m << 13);
  else bits = s | ((e + 112) << 23) | (m << 13);
  float f; std::memcpy(&f, &bits, 4); return f;
}

inline uint16_t float_to_half(float f) {  // round-to-nearest-even
  uint32_t x; std::memcpy(&x, &f, 4);
  uint32_t s = (x >> 16) & 0x8000u, a = x & 0x7FFFFFFFu;
  if (a >= 0x7F800000u) return (uint16_t)(s | 0x7C00u | ((a > 0x7F800000u) ? 0x200u : 0u));
  if (a >= 0x47800000u) {  // >= 65536 -> may still round to max; handle exactly
    return (uint16_t)(s | 0x7C00u);
  }
  if (a < 0x38800000u) {  // subnormal half or zero
    if (a < 0x33000000u) return (uint16_t)s;  // < 2^-25 -> 0
    uint32_t e = a >> 23, m = (a & 0x7FFFFFu) | 0x800000u;
    uint32_t shift = 126 - e;  // 14..24
    uint32_t r = m >> shift, rem = m & ((1u << shift) - 1), half = 1u << (shift - 1);
    if (rem > half || (rem == half && (r & 1))) r++;
    return (uint16_t)(s | r);
  }
  uint32_t e = (a >> 23) - 112, m = a & 0x7FFFFFu;
  uint32_t r = (e << 10) | (m >> 13), rem = m & 0x1FFFu;
  if (rem > 0x1000u || (rem == 0x1000u && (r & 1))) r++;  // may carry into exponent -> inf, correct
  return (uint16_t)(s | r);
}

inline float unorm_store(float v, float maxv) {  // returns the integer code as float
  if (!(v > 0.0f)) return 0.0f;  // also NaN
  if (v > 1.0f) v = 1.0f;
  return std::floor(v * maxv + 0.5f);
}

const float* srgb_decode_table();  // 256 entries, built once in double precision
float srgb_encode(float linear);   // returns value in [0,1] before the unorm8 store

// ---- texel access --------------------------------------------------------
inline const uint8_t* texel_ptr(const Image& im, int x, int y, int lod) {
  return (const uint8_t*)im.base + im.mip_offset[lod] + (size_t)y * im.row_pitch[lod] + (size_t)x * texel_size(im.format);
}

inline vec4 decode_texel(uint32_t fmt, const uint8_t* p) {
  switch (fmt) {
    case FMT_D24_UNORM_S8_UINT: { uint32_t k; std::memcpy(&k, p, 4); return vec4((float)(k & 0xFFFFFFu) / 16777215.0f, 0, 0, 1); }
    case FMT_R8_UNORM: return vec4(p[0] / 255.0f, 0, 0, 1);
    case FMT_R8G8B8A8_UNORM: return vec4(p[0] / 255.0f, p[1] / 255.0f, p[2] / 255.0f, p[3] / 255.0f);
    case FMT_R8G8B8A8_SRGB: { const float* t = srgb_decode_table(); return vec4(t[p[0]], t[p[1]], t[p[2]], p[3] / 255.0f); }
    case FMT_R16_UNORM: { uint16_t a; std::memcpy(&a, p, 2); return vec4(a / 65535.0f, 0, 0, 1); }
    case FMT_R16G16_UNORM: { uint16_t a[2]; std::memcpy(a, p, 4); return vec4(a[0] / 65535.0f, a[1] / 65535.0f, 0, 1); }
    case FMT_R16G16B16A16_UNORM: { uint16_t a[4]; std::memcpy(a, p, 8); return vec4(a[0] / 65535.0f, a[1] / 65535.0f, a[2] / 65535.0f, a[3] / 65535.0f); }
    case FMT_R16_SFLOAT: { uint16_t a; std::memcpy(&a, p, 2); return vec4(half_to_float(a), 0, 0, 1); }
    case FMT_R16G16_SFLOAT: { uint16_t a[2]; std::memcpy(a, p, 4); return vec4(half_to_float(a[0]), half_to_float(a[1]), 0, 1); }
    case FMT_R16G16B16A16_SFLOAT: { uint16_t a[4]; std::memcpy(a, p, 8); return vec4(half_to_float(a[0]), half_to_float(a[1]), half_to_float(a[2]), half_to_float(a[3])); }
    case FMT_R32_SFLOAT: { float a; std::memcpy(&a, p, 4); return vec4(a, 0, 0, 1); }
    case FMT_R32G32B32A32_SFLOAT: { float a[4]; std::memcpy(a, p, 16); return vec4(a[0], a[1], a[2], a[3]); }
  }
  return vec4();
}

// view = (image, base_mip): lod below is relative to the view, like a VkImageView.
struct View {
  const Image* im; int base_mip; int mips;
  View(const Image& i, int base = 0, int count = -1) : im(&i), base_mip(base) {
    int avail = (int)i.mip_count - base; mips = (count < 0 || count > avail) ? avail : count;
  }
  int w(int lod = 0) const { return mip_dim(im->width, base_mip + lod); }
  int h(int lod = 0) const { return mip_dim(im->height, base_mip + lod); }
};

inline ivec2 textureSize(const View& v, int lod) { return ivec2(v.w(lod), v.h(lod)); }

inline vec4 texelFetch(const View& v, ivec2 p, int lod) {
  if (lod < 0 || lod >= v.mips) return vec4();
  int w = v.w(lod), h = v.h(lod);
  if (p.x < 0 || p.y < 0 || p.x >= w || p.y >= h) return vec4();
  return decode_texel(v.im->format, texel_ptr(*v.im, p.x, p.y, v.base_mip + lod));
}

inline vec4 textureLodOffset(const View& v, vec2 uv, int lod, ivec2 off) {
  if (lod >= v.mips) lod = v.mips - 1;  // sampler clamps lod to the view
  int w = v.w(lod), h = v.h(lod);
  float x = uv.x * (float)w - 0.5f, y = uv.y * (float)h - 0.5f;
  float fx0 = std::floor(x), fy0 = std::floor(y);
  float wx = x - fx0, wy = y - fy0;
  int x0 = (int)fx0 + off.x, y0 = (int)fy0 + off.y, x1 = x0 + 1, y1 = y0 + 1;
  x0 = x0 < 0 ? 0 : (x0 > w - 1 ? w - 1 : x0); x1 = x1 < 0 ? 0 : (x1 > w - 1 ? w - 1 : x1);
  y0 = y0 < 0 ? 0 : (y0 > h - 1 ? h - 1 : y0); y1 = y1 < 0 ? 0 : (y1 > h - 1 ? h - 1 : y1);
  int L = v.base_mip + lod; uint32_t f = v.im->format;
  vec4 t00 = decode_texel(f, texel_ptr(*v.im, x0, y0, L)), t10 = decode_texel(f, texel_ptr(*v.im, x1, y0, L));
  vec4 t01 = decode_texel(f, texel_ptr(*v.im, x0, y1, L)), t11 = decode_texel(f, texel_ptr(*v.im, x1, y1, L));
  return mix(mix(t00, t10, wx), mix(t01, t11, wx), wy);
}
inline vec4 textureLod(const View& v, vec2 uv, int lod) { return textureLodOffset(v, uv, lod, ivec2(0, 0)); }
inline vec4 texture(const View& v, vec2 uv) { return textureLodOffset(v, uv, 0, ivec2(0, 0)); }
inline vec4 textureOffset(const View& v, vec2 uv, ivec2 off) { return textureLodOffset(v, uv, 0, off); }

// ---- stores --------------------------------------------------------------
inline void imageStore(const Image& im, int lod, ivec2 p, vec4 c) {
  int w = mip_dim(im.width, lod), h = mip_dim(im.height, lod);
  if (p.x < 0 || p.y < 0 || p.x >= w || p.y >= h) return;
  uint8_t* q = (uint8_t*)texel_ptr(im, p.x, p.y, lod);
  float v[4] = {c.x, c.y, c.z, c.w};
  switch (im.format) {
    case FMT_R8_UNORM: q[0] = (uint8_t)unorm_store(v[0], 255.0f); break;
    case FMT_R8G8B8A8_UNORM: for (int i = 0; i < 4; i++) q[i] = (uint8_t)unorm_store(v[i], 255.0f); break;
    case FMT_R8G8B8A8_SRGB: for (int i = 0; i < 3; i++) q[i] = (uint8_t)unorm_store(srgb_encode(v[i]), 255.0f); q[3] = (uint8_t)unorm_store(v[3], 255.0f); break;
    case FMT_R16G16_UNORM: case FMT_R16G16B16A16_UNORM: case FMT_R16_UNORM: {
      int n = im.format == FMT_R16_UNORM ? 1 : (im.format == FMT_R16G16_UNORM ? 2 : 4);
      for (int i = 0; i < n; i++) { uint16_t u = (uint16_t)unorm_store(v[i], 65535.0f); std::memcpy(q + 2 * i, &u, 2); } break; }
    case FMT_R16_SFLOAT: case FMT_R16G16_SFLOAT: case FMT_R16G16B16A16_SFLOAT: {
      int n = im.format == FMT_R16_SFLOAT ? 1 : (im.format == FMT_R16G16_SFLOAT ? 2 : 4);
      for (int i = 0; i < n; i++) { uint16_t u = float_to_half(v[i]); std::memcpy(q + 2 * i, &u, 2); } break; }
    case FMT_R32_SFLOAT: std::memcpy(q, v, 4); break;
    case FMT_R32G32B32A32_SFLOAT: std::memcpy(q, v, 16); break;
    case FMT_D24_UNORM_S8_UINT: { uint32_t k = (uint32_t)unorm_store(v[0], 16777215.0f); std::memcpy(q, &k, 4); break; }
  }
}
inline vec4 imageLoad(const Image& im, int lod, ivec2 p) { return texelFetch(View(im, lod, 1), p, 0); }

}  // namespace orc
