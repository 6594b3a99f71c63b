// common.cuh — shared device/host helpers of the sm_100a image-space passes.
//
// Texel codecs and the sampler follow the semantics pinned in SURVEY.md App. A
// (reference src/gpu/samplers.hpp:36-55 DEFAULT_SAMPLER = LINEAR / CLAMP_TO_EDGE;
// D24 texel layout reference src/main.cpp:139):
//   bilinear: x = u*W - 0.5, i0 = floor(x), w = x - i0, clamp indices to the edge,
//   sRGB decoded before weighting, mix in x then y; texelFetch out of range = 0.
// Coordinate arithmetic that selects texels uses the non-contracting intrinsics
// (__fmul_rn/__fsub_rn) so the discrete choice is identical to an unfused evaluation.
#pragma once
#include <cuda_runtime.h>
#ifdef __CUDACC__
#include "f32x2.cuh"
#endif
#include <cuda_fp16.h>
#include <stdint.h>
#include "../../include/vkr_passes.h"

namespace vkr {

// ---- host side ---------------------------------------------------------------
void set_error(const char* fmt, ...);
int check_launch(const char* what);
const float* srgb_decode_table_device();  // 256 floats on the current device (built once per device)

// One mip level of an image as the kernels see it.
struct Mip {
  const unsigned char* p;
  int w, h;
  unsigned pitch;
  unsigned long long ls;   // bytes between array layers (0 = one image shared by every layer)
};

inline int mip_dim(uint32_t d, uint32_t lod) { uint32_t v = d >> lod; return v ? (int)v : 1; }

// Why an image was refused, kept for the next set_error() on this thread (appended to the caller's message).
void note_image_problem(const char* fmt, ...);

// The kernels read rows as 128-bit (64-bit for RGBA16) vectors and index rows by pitch: base, every mip offset and every
// row pitch must be 16-byte aligned and a row must hold its texels.  vkr_image_layout() produces exactly that; an image
// described by hand (vkr_app_bind_gbuffer, a direct C-ABI caller) is refused here instead of faulting on the device.
inline bool image_layout_ok(const VkrImage* im) {
  const uint32_t bpp = vkr_format_texel_size(im->format);
  if (!bpp) { note_image_problem("unknown format %u", im->format); return false; }
  if ((uintptr_t)im->base & 15u) { note_image_problem("base %p is not 16-byte aligned", im->base); return false; }
  for (uint32_t i = 0; i < im->mip_count; i++) {
    const uint32_t w = mip_dim(im->width, i);
    if (im->mip_offset[i] & 15u) { note_image_problem("mip_offset[%u] = %llu is not 16-byte aligned", i, (unsigned long long)im->mip_offset[i]); return false; }
    if (im->row_pitch[i] & 15u) { note_image_problem("row_pitch[%u] = %u is not a multiple of 16", i, im->row_pitch[i]); return false; }
    if (im->row_pitch[i] < w * bpp) { note_image_problem("row_pitch[%u] = %u is smaller than a row of %u texels x %u bytes", i, im->row_pitch[i], w, bpp); return false; }
  }
  return true;
}

inline bool image_ok(const VkrImage* im, uint32_t format) {
  return im && im->base && im->format == format && im->mip_count >= 1 && im->mip_count <= VKR_MAX_MIPS && im->width && im->height &&
         image_layout_ok(im);
}

inline Mip make_mip(const VkrImage* im, uint32_t lod) {
  Mip m;
  m.p = (const unsigned char*)im->base + im->mip_offset[lod];
  m.w = mip_dim(im->width, lod);
  m.h = mip_dim(im->height, lod);
  m.pitch = im->row_pitch[lod];
  m.ls = 0;
  return m;
}

// ---- extended launches (VkrLaunchExt): array layers on grid.z, row windows as a block-row offset --------------
struct LaunchX {
  int yb0;   // first block row of the window (blockIdx.y is relative to it)
};
struct PeerOut {            // fused gather: extra destinations of a pass's output stores (VkrLaunchExt.peer_delta)
  int n;
  long long d[7];
};
struct ExtInfo {
  int layers = 1;
  int yb0 = 0, yblocks = 0;
  const VkrLaunchExt* e = nullptr;
  uint64_t stride(int i) const { return (e && e->layers > 1) ? e->layer_stride[i] : 0; }
  void* scratch() const { return e ? e->params_scratch : nullptr; }
  unsigned first() const { return e ? e->first_param : 0u; }   // block of the prepared set the launch's layer 0 uses
  bool prepared() const { return e && (e->flags & VKR_EXT_PARAMS_PREPARED); }
  bool prepare_only() const { return e && (e->flags & VKR_EXT_PREPARE_ONLY); }
  PeerOut peers() const { PeerOut p; p.n = e ? (int)e->peer_count : 0; for (int i = 0; i < 7; i++) p.d[i] = (e && i < p.n) ? (long long)e->peer_delta[i] : 0; return p; }
};
// Validates ext against a pass whose output has out_h rows processed in tiles of tile_h rows.
int resolve_ext(const VkrLaunchExt* e, uint32_t out_h, int tile_h, ExtInfo* o, const char* what, bool peers_ok = false);
inline Mip make_mip(const VkrImage* im, uint32_t lod, const ExtInfo& x, int arg_index) { Mip m = make_mip(im, lod); m.ls = x.stride(arg_index); return m; }
// Per-layer kernel parameter blocks: converts nothing, just ships `count` POD structs to the device scratch on `stream`.
int upload_params(const void* host, size_t elem, int count, void* scratch, cudaStream_t st, const char* what);

struct CamK {  // camera constants; tg = tanf(fovy/2) evaluated once on the host
  float aspect, znear, zfar, tg;
};
CamK make_cam(float fovy, float aspect, float znear, float zfar);

struct Dim { float w, h, rw, rh; };   // extent and RN(1/extent): pixel-centre uv without a division (see pixel_uv)
inline Dim make_dim(uint32_t w, uint32_t h) { Dim d; d.w = (float)w; d.h = (float)h; d.rw = 1.0f / d.w; d.rh = 1.0f / d.h; return d; }
inline bool same_extent(const VkrImage* a, const VkrImage* b) { return a->width == b->width && a->height == b->height; }

struct Mat4 { float m[16]; };  // column-major
inline Mat4 make_mat(const float* p) { Mat4 r; for (int i = 0; i < 16; i++) r.m[i] = p[i]; return r; }

#ifdef __CUDACC__
// ---- device side -------------------------------------------------------------
__device__ __forceinline__ Mip layer_of(Mip m, int z) { m.p += (size_t)z * m.ls; return m; }
#define VKR_BLOCK_Y ((int)blockIdx.y + lx.yb0)
#define VKR_PI 3.1415926535897932384626433832795f

// Unfused fp32: the oracle evaluates every expression as separate IEEE mul/add/sub (it is
// built with -ffp-contract=off).  The geometry helpers below use the non-contracting
// intrinsics so that every DISCRETE decision derived from them (texel selection, history
// rejection, closest-depth pick, blur radius, thickness break) is taken on the same bits as
// the oracle; colour accumulation in the kernels is left to the compiler (FMA allowed).
#define FM(a, b) __fmul_rn((a), (b))
#define FA(a, b) __fadd_rn((a), (b))
#define FS(a, b) __fsub_rn((a), (b))

// GLSL-spec min/max (NaN behaviour differs from fminf/fmaxf).
__device__ __forceinline__ float gmin(float x, float y) { return (y < x) ? y : x; }
__device__ __forceinline__ float gmax(float x, float y) { return (x < y) ? y : x; }
__device__ __forceinline__ float gclamp(float x, float lo, float hi) { return gmin(gmax(x, lo), hi); }
__device__ __forceinline__ float gmix(float a, float b, float t) { return FA(FM(a, FS(1.0f, t)), FM(b, t)); }

struct f3 { float x, y, z; };
__device__ __forceinline__ f3 mk3(float x, float y, float z) { f3 r; r.x = x; r.y = y; r.z = z; return r; }
__device__ __forceinline__ f3 operator+(f3 a, f3 b) { return mk3(FA(a.x, b.x), FA(a.y, b.y), FA(a.z, b.z)); }
__device__ __forceinline__ f3 operator-(f3 a, f3 b) { return mk3(FS(a.x, b.x), FS(a.y, b.y), FS(a.z, b.z)); }
__device__ __forceinline__ f3 operator-(f3 a) { return mk3(-a.x, -a.y, -a.z); }
__device__ __forceinline__ f3 operator*(f3 a, float s) { return mk3(FM(a.x, s), FM(a.y, s), FM(a.z, s)); }
__device__ __forceinline__ f3 operator*(float s, f3 a) { return mk3(FM(s, a.x), FM(s, a.y), FM(s, a.z)); }
__device__ __forceinline__ f3 operator*(f3 a, f3 b) { return mk3(FM(a.x, b.x), FM(a.y, b.y), FM(a.z, b.z)); }
__device__ __forceinline__ f3 operator/(f3 a, float s) { return mk3(a.x / s, a.y / s, a.z / s); }
__device__ __forceinline__ float dot3(f3 a, f3 b) { return FA(FA(FM(a.x, b.x), FM(a.y, b.y)), FM(a.z, b.z)); }
__device__ __forceinline__ float len3(f3 a) { return sqrtf(dot3(a, a)); }
__device__ __forceinline__ f3 norm3(f3 a) { float s = 1.0f / sqrtf(dot3(a, a)); return a * s; }
__device__ __forceinline__ f3 cross3(f3 a, f3 b) {
  return mk3(FS(FM(a.y, b.z), FM(b.y, a.z)), FS(FM(a.z, b.x), FM(b.z, a.x)), FS(FM(a.x, b.y), FM(b.x, a.y)));
}
__device__ __forceinline__ f3 mix3(f3 a, f3 b, float t) { return a * FS(1.0f, t) + b * t; }
__device__ __forceinline__ f3 reflect3(f3 I, f3 N) { return I - FM(2.0f, dot3(N, I)) * N; }

// M * (v, w)  — column-major, left-to-right accumulation like the oracle.
__device__ __forceinline__ float m4row(const Mat4& M, int r, f3 v, float w) {
  return FA(FA(FA(FM(M.m[r], v.x), FM(M.m[4 + r], v.y)), FM(M.m[8 + r], v.z)), FM(M.m[12 + r], w));
}
__device__ __forceinline__ f3 mul_m4(const Mat4& M, f3 v, float w) { return mk3(m4row(M, 0, v, w), m4row(M, 1, v, w), m4row(M, 2, v, w)); }
__device__ __forceinline__ float4 mul_m4_4(const Mat4& M, f3 v, float w) {
  return make_float4(m4row(M, 0, v, w), m4row(M, 1, v, w), m4row(M, 2, v, w), m4row(M, 3, v, w));
}

// ---- shared shader math (reference src/shaders/include/gbuffer_encode.glsl) -----
__device__ __forceinline__ float sign_nz(float k) { return (k >= 0.0f) ? 1.0f : -1.0f; }

__device__ __forceinline__ f3 decode_normal(float u, float v) {           // gbuffer_encode.glsl:29-37
  u = FS(FM(2.f, u), 1.f); v = FS(FM(2.f, v), 1.f);
  f3 n = mk3(u, v, FS(FS(1.0f, fabsf(u)), fabsf(v)));
  if (n.z < 0.0f) {
    float tx = FM(FS(1.0f, fabsf(n.y)), sign_nz(n.x));
    float ty = FM(FS(1.0f, fabsf(n.x)), sign_nz(n.y));
    n.x = tx; n.y = ty;
  }
  return norm3(n);
}

__device__ __forceinline__ float linearize_depth2(float d, float n, float f) {   // :53-56
  return FM(n, f) / FS(FM(d, FS(f, n)), f);
}

__device__ __forceinline__ f3 reconstruct_view_vec(float u, float v, float d, const CamK& c) {  // :58-69
  float z = linearize_depth2(d, c.znear, c.zfar);
  float xd = FS(FM(2.f, u), 1.f);
  float yd = FS(FM(2.f, v), 1.f);
  float x = FM(-(xd), FM(FM(z, c.aspect), c.tg));
  float y = FM(-(yd), FM(z, c.tg));
  return mk3(x, y, z);
}

__device__ __forceinline__ f3 project_view_vec(f3 v, const CamK& c) {                          // :75-84
  float n = c.znear, f = c.zfar;
  float z = v.z;
  float depth = FA(f / FS(f, n), FM(f, n) / FM(z, FS(f, n)));
  float pu = v.x / FM(FM(-v.z, c.tg), c.aspect);
  float pv = v.y / FM(-z, c.tg);
  return mk3(FA(FM(0.5f, pu), 0.5f), FA(FM(0.5f, pv), 0.5f), depth);
}

// ---- texel decode ----------------------------------------------------------------
// Correctly rounded k / D for an integer-valued k and a constant D in 3 instructions instead of
// the ~12 of the generic IEEE division (MUFU.RCP + Newton + FCHK + slow-path branch): with
// r = RN(1/D), q = RN(k*r), the FMA residual e = k - q*D is exact and RN(q + e*r) is the correctly
// rounded quotient (Markstein).  vkr_selftest_decode() checks all 2^24 + 2^16 + 2^8 inputs against
// __fdiv_rn on the device (tests/test_parity_gpu.py::test_unorm_decode_is_exact).
__device__ __forceinline__ float div_by_const(float k, float d, float r) {
  float q = __fmul_rn(k, r);
  float e = __fmaf_rn(-q, d, k);
  return __fmaf_rn(e, r, q);
}
__device__ __forceinline__ float d24_to_float(uint32_t k) { return div_by_const((float)(k & 0xFFFFFFu), 16777215.0f, 1.0f / 16777215.0f); }
__device__ __forceinline__ float unorm16_to_float(uint32_t k) { return div_by_const((float)k, 65535.0f, 1.0f / 65535.0f); }
__device__ __forceinline__ float unorm8_to_float(uint32_t k) { return div_by_const((float)k, 255.0f, 1.0f / 255.0f); }

__device__ __forceinline__ uint32_t ld32(const Mip& m, int x, int y) {
  return __ldg((const uint32_t*)(m.p + (size_t)y * m.pitch) + x);
}
__device__ __forceinline__ uint2 ld64(const Mip& m, int x, int y) {
  return __ldg((const uint2*)(m.p + (size_t)y * m.pitch) + x);
}
__device__ __forceinline__ uint16_t ld16(const Mip& m, int x, int y) {
  return __ldg((const uint16_t*)(m.p + (size_t)y * m.pitch) + x);
}
__device__ __forceinline__ bool in_range(const Mip& m, int x, int y) { return x >= 0 && y >= 0 && x < m.w && y < m.h; }

// texelFetch of a D24 texel as float; out of range = 0.
__device__ __forceinline__ float fetch_d24(const Mip& m, int x, int y) {
  return in_range(m, x, y) ? d24_to_float(ld32(m, x, y)) : 0.0f;
}

// Bilinear footprint: indices (clamped) and weights for uv on a w x h level.
struct Tap { int x0, x1, y0, y1; float wx, wy; };
__device__ __forceinline__ Tap make_tap(float u, float v, int w, int h, int ox = 0, int oy = 0) {
  float x = __fsub_rn(__fmul_rn(u, (float)w), 0.5f), y = __fsub_rn(__fmul_rn(v, (float)h), 0.5f);
  float fx = floorf(x), fy = floorf(y);
  Tap t;
  t.wx = __fsub_rn(x, fx); t.wy = __fsub_rn(y, fy);
  int x0 = (int)fx + ox, y0 = (int)fy + oy;
  t.x0 = min(max(x0, 0), w - 1); t.x1 = min(max(x0 + 1, 0), w - 1);
  t.y0 = min(max(y0, 0), h - 1); t.y1 = min(max(y0 + 1, 0), h - 1);
  return t;
}
__device__ __forceinline__ float bilerp(float t00, float t10, float t01, float t11, float wx, float wy) {
  return gmix(gmix(t00, t10, wx), gmix(t01, t11, wx), wy);
}

// Two D24 texels decoded at once: div_by_const on a packed pair (same three roundings per lane).
__device__ __forceinline__ f2 d24x2_to_float(uint32_t ka, uint32_t kb) {
  const f2 k = pk((float)(ka & 0xFFFFFFu), (float)(kb & 0xFFFFFFu)), r = dup(1.0f / 16777215.0f);
  const f2 q = mul2(k, r);
  return fma2(fma2(q, dup(-16777215.0f), k), r, q);
}
// textureLod of a D24 level, exact: make_tap + bilerp above, evaluated on packed fp32 (x with y, the left texel column
// with the right one).  Every lane rounds like the scalar unfused operation it replaces, so the result is bit-identical
// to bilerp(d24_to_float(..) x 4, wx, wy); it just takes ~13 fewer issue slots.
__device__ __forceinline__ float sample_d24(const Mip& m, float u, float v, int ox = 0, int oy = 0) {
  const f2 one = one2();
  const f2 xy = mul_add2_unfused(pk(u, v), pk((float)m.w, (float)m.h), dup(-0.5f), one);
  float x, y, wx, wy, omx, omy, top, bot;
  unpk(xy, x, y);
  const float fx = floorf(x), fy = floorf(y);
  const f2 w2 = sub2(xy, pk(fx, fy));
  unpk(w2, wx, wy);
  unpk(sub2(dup(1.0f), w2), omx, omy);
  const int x0 = (int)fx + ox, y0 = (int)fy + oy;
  const int xa = min(max(x0, 0), m.w - 1), xb = min(max(x0 + 1, 0), m.w - 1), ya = min(max(y0, 0), m.h - 1), yb = min(max(y0 + 1, 0), m.h - 1);
  const f2 left = d24x2_to_float(ld32(m, xa, ya), ld32(m, xa, yb)), right = d24x2_to_float(ld32(m, xb, ya), ld32(m, xb, yb));
  unpk(gmix2(left, right, omx, wx, one), top, bot);
  return FA(FM(top, omy), FM(bot, wy));
}

__device__ __forceinline__ float2 unpack_rg16unorm(uint32_t k) { return make_float2(unorm16_to_float(k & 0xFFFFu), unorm16_to_float(k >> 16)); }
__device__ __forceinline__ float2 unpack_rg16f(uint32_t k) { return __half22float2(*reinterpret_cast<const __half2*>(&k)); }

__device__ __forceinline__ float2 sample_rg16unorm(const Mip& m, float u, float v) {
  Tap t = make_tap(u, v, m.w, m.h);
  float2 a = unpack_rg16unorm(ld32(m, t.x0, t.y0)), b = unpack_rg16unorm(ld32(m, t.x1, t.y0));
  float2 c = unpack_rg16unorm(ld32(m, t.x0, t.y1)), d = unpack_rg16unorm(ld32(m, t.x1, t.y1));
  return make_float2(bilerp(a.x, b.x, c.x, d.x, t.wx, t.wy), bilerp(a.y, b.y, c.y, d.y, t.wx, t.wy));
}
__device__ __forceinline__ f3 sample_gbuffer_normal(const Mip& m, float u, float v) {   // gbuffer_encode.glsl:39-43
  float2 e = sample_rg16unorm(m, u, v);
  return decode_normal(e.x, e.y);
}
__device__ __forceinline__ float2 sample_rg16f(const Mip& m, float u, float v, int ox = 0, int oy = 0) {
  Tap t = make_tap(u, v, m.w, m.h, ox, oy);
  float2 a = unpack_rg16f(ld32(m, t.x0, t.y0)), b = unpack_rg16f(ld32(m, t.x1, t.y0));
  float2 c = unpack_rg16f(ld32(m, t.x0, t.y1)), d = unpack_rg16f(ld32(m, t.x1, t.y1));
  return make_float2(bilerp(a.x, b.x, c.x, d.x, t.wx, t.wy), bilerp(a.y, b.y, c.y, d.y, t.wx, t.wy));
}
__device__ __forceinline__ float sample_r32f(const Mip& m, float u, float v) {
  Tap t = make_tap(u, v, m.w, m.h);
  return bilerp(__uint_as_float(ld32(m, t.x0, t.y0)), __uint_as_float(ld32(m, t.x1, t.y0)),
                __uint_as_float(ld32(m, t.x0, t.y1)), __uint_as_float(ld32(m, t.x1, t.y1)), t.wx, t.wy);
}

// sRGB8 texel -> linear rgb (+ linear alpha in .w); srgb = 256-entry decode table.
__device__ __forceinline__ float4 unpack_srgb8(uint32_t k, const float* __restrict__ srgb) {
  return make_float4(__ldg(srgb + (k & 255u)), __ldg(srgb + ((k >> 8) & 255u)), __ldg(srgb + ((k >> 16) & 255u)), unorm8_to_float(k >> 24));
}
__device__ __forceinline__ float4 sample_srgb8(const Mip& m, float u, float v, const float* __restrict__ srgb) {
  Tap t = make_tap(u, v, m.w, m.h);
  float4 a = unpack_srgb8(ld32(m, t.x0, t.y0), srgb), b = unpack_srgb8(ld32(m, t.x1, t.y0), srgb);
  float4 c = unpack_srgb8(ld32(m, t.x0, t.y1), srgb), d = unpack_srgb8(ld32(m, t.x1, t.y1), srgb);
  return make_float4(bilerp(a.x, b.x, c.x, d.x, t.wx, t.wy), bilerp(a.y, b.y, c.y, d.y, t.wx, t.wy),
                     bilerp(a.z, b.z, c.z, d.z, t.wx, t.wy), bilerp(a.w, b.w, c.w, d.w, t.wx, t.wy));
}
__device__ __forceinline__ f3 unpack_rgb8unorm(uint32_t k) {
  return mk3(unorm8_to_float(k & 255u), unorm8_to_float((k >> 8) & 255u), unorm8_to_float((k >> 16) & 255u));
}

// ---- value-only fast paths -------------------------------------------------------
// Precision policy: a quantity needs the unfused, IEEE-divided evaluation above only if a DISCONTINUOUS
// choice hangs on it (a texelFetch cell index, an argmin between near-tied candidates, a radius floor).
// Bilinear results are continuous in their coordinates and thresholds on continuous quantities flip on a
// ~1e-7 fraction of pixels, so everything below may use FMA contraction, weight-form bilinear (4 FMA per
// channel instead of 12 unfused ops), MUFU reciprocal / rsqrt.  Deviation from the oracle stays ~1e-6
// relative, three orders of magnitude inside the stated output tolerances (tests/parity.py).

// (px + 0.5) / W and px / W, correctly rounded in 3 instructions (checked exhaustively for every
// W <= 16384, px < W against the IEEE division: tests/test_abi.py::test_pixel_uv_division_is_exact).
__device__ __forceinline__ float pixel_uv(int p, float dim, float rdim) { return div_by_const((float)p + 0.5f, dim, rdim); }
__device__ __forceinline__ float texel_uv(int p, float dim, float rdim) { return div_by_const((float)p, dim, rdim); }

struct TapW { int x0, x1, y0, y1; float w00, w10, w01, w11; };
__device__ __forceinline__ TapW make_tapw(const Tap& t) {
  TapW r;
  r.x0 = t.x0; r.x1 = t.x1; r.y0 = t.y0; r.y1 = t.y1;
  r.w11 = t.wx * t.wy; r.w10 = t.wx - r.w11; r.w01 = t.wy - r.w11; r.w00 = (1.0f - t.wx) - r.w01;
  return r;
}
__device__ __forceinline__ TapW make_tapw(float u, float v, int w, int h, int ox = 0, int oy = 0) { return make_tapw(make_tap(u, v, w, h, ox, oy)); }
__device__ __forceinline__ float wsum(float a, float b, float c, float d, const TapW& w) {
  return fmaf(d, w.w11, fmaf(c, w.w01, fmaf(b, w.w10, a * w.w00)));
}
// A footprint whose weights are within 2^-10 of a corner (a same-size image sampled at its own pixel
// centres: |w - {0,1}| <= px * 2^-23) collapses to that texel; the dropped term is < 2.5e-4 of the
// neighbour difference.  Only used for colour-like values, never for depth.
__device__ __forceinline__ bool tap_is_texel(const Tap& t, int& x, int& y) {
  const float e = 0.0009765625f;
  bool okx = (t.wx < e) || (t.wx > 1.0f - e), oky = (t.wy < e) || (t.wy > 1.0f - e);
  x = (t.wx < 0.5f) ? t.x0 : t.x1;
  y = (t.wy < 0.5f) ? t.y0 : t.y1;
  return okx && oky;
}

// single-MUFU reciprocal / square root / division (2 ulp): the IEEE forms cost 8-12 instructions plus a slow-path call
__device__ __forceinline__ float rcp_fast(float x) { float r; asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x)); return r; }
__device__ __forceinline__ float sqrt_fast(float x) { float r; asm("sqrt.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x)); return r; }
__device__ __forceinline__ float rsqrt_fast(float x) { float r; asm("rsqrt.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x)); return r; }
__device__ __forceinline__ float div_fast(float a, float b) { return a * rcp_fast(b); }
__device__ __forceinline__ float dot3f(f3 a, f3 b) { return fmaf(a.z, b.z, fmaf(a.y, b.y, a.x * b.x)); }
__device__ __forceinline__ f3 norm3f(f3 a) { float s = rsqrtf(dot3f(a, a)); return mk3(a.x * s, a.y * s, a.z * s); }
__device__ __forceinline__ float len3f(f3 a) { return sqrt_fast(dot3f(a, a)); }
__device__ __forceinline__ f3 sub3f(f3 a, f3 b) { return mk3(a.x - b.x, a.y - b.y, a.z - b.z); }
__device__ __forceinline__ f3 mul_m4f(const Mat4& M, f3 v) {   // M * (v, 1), xyz
  return mk3(fmaf(M.m[8], v.z, fmaf(M.m[4], v.y, fmaf(M.m[0], v.x, M.m[12]))), fmaf(M.m[9], v.z, fmaf(M.m[5], v.y, fmaf(M.m[1], v.x, M.m[13]))),
             fmaf(M.m[10], v.z, fmaf(M.m[6], v.y, fmaf(M.m[2], v.x, M.m[14]))));
}
__device__ __forceinline__ f3 mul_m3f(const Mat4& M, f3 v) {   // M * (v, 0), xyz
  return mk3(fmaf(M.m[8], v.z, fmaf(M.m[4], v.y, M.m[0] * v.x)), fmaf(M.m[9], v.z, fmaf(M.m[5], v.y, M.m[1] * v.x)),
             fmaf(M.m[10], v.z, fmaf(M.m[6], v.y, M.m[2] * v.x)));
}
// a / b rounded to nearest like the IEEE division for normal-range operands (reciprocal refined once, quotient corrected
// by its exact FMA residual) without the compiler's FCHK + slow-path call: 6 instructions.  Used where a quotient feeds
// a threshold through an ill-conditioned expression and must carry the oracle's bits.
__device__ __forceinline__ float div_rn_normal(float a, float b) {
  float r = rcp_fast(b);
  r = __fmaf_rn(__fmaf_rn(-b, r, 1.0f), r, r);
  float q = __fmul_rn(a, r);
  return __fmaf_rn(__fmaf_rn(-q, b, a), r, q);
}
struct CamF { float aspect_tg, tg, nf, fmn, zfar; };   // hoisted products for the fast reconstruct
__device__ __forceinline__ CamF make_camf(const CamK& c) {
  CamF r; r.aspect_tg = c.aspect * c.tg; r.tg = c.tg; r.nf = FM(c.znear, c.zfar); r.fmn = FS(c.zfar, c.znear); r.zfar = c.zfar; return r;
}
// d*(f-n) - f cancels catastrophically for far depths (|den| ~ 0.1 against terms of 80): a fused or reordered
// evaluation moves z by up to ~1e-2 view units there, which is the size of the reprojection thresholds (0.01 .. 0.2)
// downstream.  So the denominator keeps the oracle's unfused rounding; only the final division is approximate (2 ulp).
__device__ __forceinline__ float linearize_fast(float d, const CamF& c) { return c.nf * rcp_fast(FS(FM(d, c.fmn), c.zfar)); }
// linearize_depth2 with the oracle's bits (exact denominator, correctly rounded quotient)
__device__ __forceinline__ float linearize_rn(float d, const CamF& c) { return div_rn_normal(c.nf, FS(FM(d, c.fmn), c.zfar)); }
__device__ __forceinline__ f3 reconstruct_fast(float u, float v, float d, const CamF& c) {
  float z = linearize_fast(d, c);
  return mk3((1.0f - 2.f * u) * (z * c.aspect_tg), (1.0f - 2.f * v) * (z * c.tg), z);
}
// project_view_vec with MUFU reciprocals (value-only); the camera-only quotients are uniform
__device__ __forceinline__ f3 project_fast(f3 v, const CamK& c) {
  const float fmn = c.zfar - c.znear;
  const float A = c.zfar / fmn, B = c.zfar * c.znear / fmn;
  const float rz = rcp_fast(v.z);
  const float itg = rcp_fast(c.tg);
  const float pu = -v.x * rz * (itg * rcp_fast(c.aspect)), pv = -v.y * rz * itg;
  return mk3(fmaf(0.5f, pu, 0.5f), fmaf(0.5f, pv, 0.5f), fmaf(B, rz, A));
}
__device__ __forceinline__ f3 decode_normal_fast(float u, float v) {
  u = fmaf(2.f, u, -1.f); v = fmaf(2.f, v, -1.f);
  f3 n = mk3(u, v, 1.0f - fabsf(u) - fabsf(v));
  if (n.z < 0.0f) {
    float tx = (1.0f - fabsf(n.y)) * sign_nz(n.x), ty = (1.0f - fabsf(n.x)) * sign_nz(n.y);
    n.x = tx; n.y = ty;
  }
  return norm3f(n);
}
__device__ __forceinline__ float2 wsum2(float2 a, float2 b, float2 c, float2 d, const TapW& w) {
  return make_float2(wsum(a.x, b.x, c.x, d.x, w), wsum(a.y, b.y, c.y, d.y, w));
}
__device__ __forceinline__ float2 sample_rg16f_fast(const Mip& m, const TapW& t) {
  return wsum2(unpack_rg16f(ld32(m, t.x0, t.y0)), unpack_rg16f(ld32(m, t.x1, t.y0)), unpack_rg16f(ld32(m, t.x0, t.y1)), unpack_rg16f(ld32(m, t.x1, t.y1)), t);
}
__device__ __forceinline__ float2 sample_rg16unorm_fast(const Mip& m, const TapW& t) {
  return wsum2(unpack_rg16unorm(ld32(m, t.x0, t.y0)), unpack_rg16unorm(ld32(m, t.x1, t.y0)), unpack_rg16unorm(ld32(m, t.x0, t.y1)),
               unpack_rg16unorm(ld32(m, t.x1, t.y1)), t);
}
// Bilinear oct-normal at uv, or the single texel when the footprint is degenerate (same-size image, own pixel centre).
__device__ __forceinline__ f3 sample_gbuffer_normal_fast(const Mip& m, float u, float v) {
  Tap t = make_tap(u, v, m.w, m.h);
  int x, y;
  float2 e = tap_is_texel(t, x, y) ? unpack_rg16unorm(ld32(m, x, y)) : sample_rg16unorm_fast(m, make_tapw(t));
  return decode_normal_fast(e.x, e.y);
}

// t / d for a run-time d in staging loops: q = umulhi(t, ceil(2^32 / d)) is the exact quotient while t < 2^16 and d < 2^16
// (two instructions instead of the ~20 of the generic 32-bit division; gtao.cu spent 5 % of its issue slots on one of those).
__device__ __forceinline__ uint32_t div_magic(uint32_t d) { return 0xFFFFFFFFu / d + 1u; }        // d = 1 wraps to 0: see div_small
__device__ __forceinline__ int div_small(int t, uint32_t magic) { return magic ? (int)__umulhi((uint32_t)t, magic) : t; }

// ---- texel encode ----------------------------------------------------------------
__device__ __forceinline__ uint32_t to_unorm(float v, float maxv) {   // clamp, NaN -> 0, floor(v*max + 0.5)
  if (!(v > 0.0f)) return 0u;
  if (v > 1.0f) v = 1.0f;
  return (uint32_t)floorf(__fadd_rn(__fmul_rn(v, maxv), 0.5f));
}
__device__ __forceinline__ uint32_t pack_rgba8unorm(float r, float g, float b, float a) {
  return to_unorm(r, 255.f) | (to_unorm(g, 255.f) << 8) | (to_unorm(b, 255.f) << 16) | (to_unorm(a, 255.f) << 24);
}
__device__ __forceinline__ uint32_t pack_half2(float a, float b) {
  __half2 h = __floats2half2_rn(a, b);
  return *reinterpret_cast<uint32_t*>(&h);
}
__device__ __forceinline__ float srgb_encode(float l) {
  if (!(l > 0.0f)) return 0.0f;
  if (l >= 1.0f) return 1.0f;
  if (l < 0.0031308f) return 12.92f * l;
  return 1.055f * powf(l, 1.0f / 2.4f) - 0.055f;
}
#endif  // __CUDACC__

}  // namespace vkr
