// ssr_trace.cu — stochastic SSR ray: GGX-VNDF direction, hierarchical Hi-Z march (<= 80
// steps, first 15 pinned to the finest level) that also gathers a horizon for AO, three
// validity tests, and the ray's (occlusion, pdf) deposited into gtao.raw for the MIS combine.
// Replaces task "SSSR_trace" (reference src/advanced_ssr.cpp:147-214,
// shaders/advanced_ssr/trace.comp:41-141,:206-268, shaders/include/screen_trace.glsl:8-45).
#include <vector>
#include "brdf.cuh"
#include "march.cuh"

#ifndef TRACE_WARP_SHAPE
#define TRACE_WARP_SHAPE 162
#endif
#ifndef TRACE_UNROLL_A
#define TRACE_UNROLL_A 1
#endif
#ifndef TRACE_V2
#define TRACE_V2 1   /* the re-stated march below (profiles/r02a_trace_variants.txt, r02c_trace_variants.txt); 0 = round 1's kernel */
#endif
namespace vkr {

struct TraceDims { float ssx, ssy; };   // level-0 extent of the pyramid view as floats

struct TraceK {
  Mat4 normal_mat;
  CamK cam;
  uint32_t frame_random;
  float max_roughness;
  CamF camf;                           // make_camf(cam) and -2 * {aspect_tg, tg}, evaluated once on the host (same single roundings):
  float m2at, m2tg;                    // constant-bank operands of the v2 march's horizon sample instead of per-step re-materialisation
};

// trace.comp:156-158, bit-defined (see oracle/passes_ssr.cpp rand_hash): unfused fp32 dot,
// sin in double rounded once to fp32, fp32 multiply, fract.
__device__ __forceinline__ float rand_hash(float u, float v) {
  float d = __fadd_rn(__fmul_rn(u, 12.9898f), __fmul_rn(v, 78.233f));
  float s = (float)sin((double)d);
  float p = __fmul_rn(s, 43758.5453f);
  return __fsub_rn(p, floorf(p));
}

template <bool LAYERED>
__global__ void __launch_bounds__(128) sssr_trace_kernel(TraceK k0, const TraceK* __restrict__ kv, LaunchX lx, PyramidView P, Mip normal, Mip material,
                                                         const float4* __restrict__ halton, Mip pdf, Mip out_rays, Mip out_occl, const float* __restrict__ srgb, TraceDims dims) {
  TraceK kl;
  unsigned long long poff = 0;                         // layer offset of the depth pyramid (one allocation per layer holds every mip)
  if (LAYERED) {
    const int z = blockIdx.z;
    kl = kv[z];
    poff = (unsigned long long)z * P.mip[0].ls;
    normal = layer_of(normal, z); material = layer_of(material, z); pdf = layer_of(pdf, z); out_rays = layer_of(out_rays, z); out_occl = layer_of(out_occl, z);
  }
  const TraceK& k = LAYERED ? kl : k0;
#if TRACE_WARP_SHAPE == 84      // a warp covers 8 x 4 pixels: neighbouring rays march alike, so a compact footprint idles fewer lanes
  const int px = blockIdx.x * 16 + ((threadIdx.x >> 5) & 1) * 8 + (threadIdx.x & 7), py = VKR_BLOCK_Y * 8 + (threadIdx.x >> 6) * 4 + ((threadIdx.x >> 3) & 3);
#elif TRACE_WARP_SHAPE == 48    // 4 x 8
  const int px = blockIdx.x * 16 + ((threadIdx.x >> 5) & 3) * 4 + (threadIdx.x & 3), py = VKR_BLOCK_Y * 8 + ((threadIdx.x >> 2) & 7);
#else                           // 16 x 2
  const int px = blockIdx.x * 16 + (threadIdx.x & 15), py = VKR_BLOCK_Y * 8 + (threadIdx.x >> 4);
#endif
  if (px >= out_rays.w || py >= out_rays.h) return;
  const float tsx = (float)out_rays.w, tsy = (float)out_rays.h;
  const float su = ((float)px + 0.5f) / tsx, sv = ((float)py + 0.5f) / tsy;                  // :44
  const CamK cam = k.cam;
  const CamF camf = make_camf(cam);

  float roughness = sample_srgb8(material, su, sv, srgb).y;                                     // :49-53
  float mg = gmix(0.0f, k.max_roughness, roughness);
  roughness = FM(mg, mg);

  Mip d0 = P.mip[0];
  d0.p += poff;
  float pixel_depth = sample_d24(d0, su, sv);
  f3 pixel_normal = norm3(mul_m4(k.normal_mat, sample_gbuffer_normal(normal, su, sv), 0.f));
  f3 view_vec = reconstruct_view_vec(su, sv, pixel_depth, cam);

  const uint32_t base_index = (uint32_t)(rand_hash(su, sv) * 128.0f);                           // :61-63
  uint32_t index = (base_index + k.frame_random) & 127u;
  float4 hv = __ldg(halton + index);
  f3 tangent = get_tangent(pixel_normal);
  f3 bitangent = norm3(cross3(pixel_normal, tangent));
  tangent = norm3(cross3(bitangent, pixel_normal));

  f3 w0 = -norm3(view_vec);
  f3 view_dir = mk3(dot3(w0, tangent), dot3(w0, bitangent), dot3(w0, pixel_normal));
  f3 brdf_norm = sampleGGXVNDF(view_dir, roughness, roughness, hv.x, hv.y);
  f3 N = brdf_norm.x * tangent + brdf_norm.y * bitangent + brdf_norm.z * pixel_normal;
  f3 R = reflect3(view_vec, N);

  f3 ray_start = project_view_vec(view_vec + 0.001f * pixel_normal, cam);                      // :79-84
  ray_start.z = FS(ray_start.z, 0.0001f);
  f3 ray_dir = project_view_vec(view_vec + R, cam) - ray_start;
  ray_dir = ray_dir * (FS(1.f, ray_start.z) / ray_dir.z);

  // ---- hierarchical_raymarch_find_hor, trace.comp:206-268 ------------------------------------
  const f3 origin = ray_start, direction = ray_dir;
  const f3 inv_direction = mk3(direction.x != 0 ? 1.0f / direction.x : MAX_T_FLOAT, direction.y != 0 ? 1.0f / direction.y : MAX_T_FLOAT,
                               direction.z != 0 ? 1.0f / direction.z : MAX_T_FLOAT);
  int current_mip = 0;
  const float ssx = (float)d0.w, ssy = (float)d0.h;
  float resx = ssx, resy = ssy;                       // get_mip_resolution(screen_size, 0)
  float rinvx = 1.0f / resx, rinvy = 1.0f / resy;
  float uox = 0.005f / ssx, uoy = 0.005f / ssy;       // 0.005 * exp2(0) / screen_size
  uox = direction.x < 0 ? -uox : uox;
  uoy = direction.y < 0 ? -uoy : uoy;
  const float fox = direction.x < 0 ? 0.f : 1.f, foy = direction.y < 0 ? 0.f : 1.f;
  float current_t;
  f3 position;
  {                                                    // initial_advance_ray, screen_trace.glsl:8-15
    float xp = FA(FM(FA(floorf(FM(resx, origin.x)), fox), rinvx), uox);
    float yp = FA(FM(FA(floorf(FM(resy, origin.y)), foy), rinvy), uoy);
    float tx = FM(FS(xp, origin.x), inv_direction.x), ty = FM(FS(yp, origin.y), inv_direction.y);
    current_t = gmin(tx, ty);
    position = origin + current_t * direction;
  }
  float h = 0.f;
  int i = 0;
  // horizon sample (:258-262): a running max of a continuous quantity -> value-only arithmetic.
  // len(v) < 0.3 <=> |v|^2 < 0.09; a zero-length v gives NaN, which both max forms ignore.
#define TRACE_HORIZON(SURFACE_Z)                                                                                                   \
  {                                                                                                                                \
    const f3 v = sub3f(reconstruct_fast(position.x, position.y, SURFACE_Z, camf), view_vec);                                       \
    const float d2 = dot3f(v, v);                                                                                                  \
    if (d2 < 0.09f) { /* most samples of a long ray are farther than 0.3: skip the normalisation */                               \
      float r;                                                                                                                     \
      asm("rsqrt.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(d2));                                                                     \
      h = fmaxf(h, dot3f(pixel_normal, v) * r);                                                                                    \
    }                                                                                                                              \
  }
  // The first 15 steps never leave mip 0 (:250), for every ray: a uniform loop with the level's extent, pointer and cell
  // size as compile-time-indexed constants, no level bookkeeping and an unconditional horizon sample.
  constexpr int unroll_a = TRACE_UNROLL_A;
#pragma unroll unroll_a
  for (; i < 15; ++i) {
    float mpx = FM(ssx, position.x), mpy = FM(ssy, position.y);
    float surface_z = fetch_pyramid(P, poff, (int)mpx, (int)mpy, 0);
    advance_ray(origin, direction, inv_direction, mpx, mpy, rinvx, rinvy, fox, foy, uox, uoy, surface_z, position, current_t);
    TRACE_HORIZON(surface_z)
  }
  while (i < 80 && current_mip >= 0) {
    float mpx = FM(resx, position.x), mpy = FM(resy, position.y);
    float surface_z = fetch_pyramid(P, poff, (int)mpx, (int)mpy, current_mip);
    bool skipped = advance_ray(origin, direction, inv_direction, mpx, mpy, rinvx, rinvy, fox, foy, uox, uoy, surface_z, position, current_t);
    current_mip += skipped ? 1 : -1;
    float f = skipped ? 0.5f : 2.0f, g = skipped ? 2.0f : 0.5f;
    resx *= f; resy *= f; rinvx *= g; rinvy *= g;
    ++i;
    if (current_mip <= 1) TRACE_HORIZON(surface_z)
  }
#undef TRACE_HORIZON
  bool valid_hit = true;                              // (i <= max) is always true, App. A.6 #8
  const f3 out_ray = position;

  if (valid_hit) {                                                                               // :94-99
    float sx = FM(fabsf(FS(out_ray.x, ray_start.x)), tsx), sy = FM(fabsf(FS(out_ray.y, ray_start.y)), tsy);
    if (gmax(sx, sy) < 2.0f) valid_hit = false;
  }
  if (valid_hit) {                                                                               // :101-109
    f3 hit_normal = mul_m4(k.normal_mat, sample_gbuffer_normal(normal, out_ray.x, out_ray.y), 0.f);
    if (dot3(hit_normal, R) > 0 || dot3(pixel_normal, R) < 0) valid_hit = false;
  }
  if (valid_hit) {                                                                               // :111-118
    float hit_depth = sample_d24(d0, out_ray.x, out_ray.y);
    float hit_z = linearize_depth2(hit_depth, cam.znear, cam.zfar);
    float ray_z = linearize_depth2(out_ray.z, cam.znear, cam.zfar);
    if (ray_z > FA(hit_z, 0.3f) || ray_z < FS(hit_z, 0.1f)) valid_hit = false;
  }
  {                                                                                              // :120-121, RGBA16_UNORM
    uint32_t lo = to_unorm(out_ray.x, 65535.f) | (to_unorm(out_ray.y, 65535.f) << 16);
    uint32_t hi = to_unorm(out_ray.z, 65535.f) | (to_unorm(valid_hit ? pixel_depth : 1.0f, 65535.f) << 16);
    *((uint2*)(out_rays.p + (size_t)py * out_rays.pitch) + px) = make_uint2(lo, hi);
  }
  {                                                                                              // :123-139
    f3 slice_normal = norm3(cross3(w0, R));
    f3 normal_projected = pixel_normal - dot3(pixel_normal, slice_normal) * slice_normal;
    f3 X = norm3(cross3(slice_normal, w0));
    float n = FS(VKR_PI / 2.0f, acosf(dot3(norm3(normal_projected), X)));
    float hh = acosf(h);
    hh = gmin(FA(n, gmin(FS(hh, n), VKR_PI / 2.0f)), hh);
    float pdfv = sampleGGXdirPDF(pdf, w0, pixel_normal, R, roughness);
    float occlusion = FM(FM(FM(1.0f / VKR_PI, len3(normal_projected)), 0.25f), gmax(FA(FA(-cosf(FS(FM(2.f, hh), n)), cosf(n)), FM(FM(2.f, hh), sinf(n))), 0.f));
    float result = isnan(occlusion) ? 0.0f : occlusion;
#ifdef VKR_TRACE_STATS
    *((uint2*)(out_occl.p + (size_t)py * out_occl.pitch) + px) = make_uint2(pack_half2(result, pdfv), pack_half2((float)i, (float)current_mip));
#else
    *((uint2*)(out_occl.p + (size_t)py * out_occl.pitch) + px) = make_uint2(pack_half2(result, pdfv), 0u);
#endif
  }
}

}  // namespace vkr

// ================================================================================================
// Version 2 of the march (default, TRACE_V2=1).  Same algorithm, re-stated for the sm_100a issue pipes:
//  * the cell index of a level-m step comes out of ONE FFMA with round-down:
//      cell * 2^m = RD(W * x + 1.5 * 2^(23+m)) - 1.5 * 2^(23+m)
//    (the magic constant makes the unit in the last place 2^m, so rounding down IS the floor; the integer index is the
//    difference of the two bit patterns).  No FRND / F2I on the quarter-rate conversion pipe and no per-level resolution
//    registers: the level is carried by the magic constant alone (level +-1 = exponent +-1).
//  * the ray is parameterised by t only; x, y, z and the horizon's (1 - 2u) * aspect * tan terms are FFMAs of t.
//  * the level table {base, pitch, extent} lives in shared memory instead of lane-indexed constant loads.
//  * the fast loops run while the ray is inside the screen minus 2^-20; leaving it (at most once per ray) falls into a
//    slow tail that restates the shader's out-of-range / truncation behaviour literally (texelFetch at trunc(), App. A.3).
// Precision policy: the step arithmetic is contracted (<= 1 rounding per uv instead of 2), the cell
// index is the floor of the exactly rounded-down product.  The depth fed to the horizon keeps the oracle's bits (exact
// D24 decode, unfused d*(f-n) - f: that difference is ill-conditioned, see linearize_fast).
// TRACE_FAST_SETUP=1 also evaluates the ray set-up and the occlusion epilogue with FMA contraction and single-MUFU
// reciprocals / square roots; the Halton hash stays bit-defined.
#ifndef TRACE_FAST_SETUP
#define TRACE_FAST_SETUP 1
#endif
#ifndef TRACE_MIN_CTAS
#define TRACE_MIN_CTAS 9                 /* 56 registers, no spill in the plain kernel: 0.3553 -> 0.3522 ms at 4K */
#endif
#ifndef TRACE_CTA_ROWS
#define TRACE_CTA_ROWS 8                /* 16 x TRACE_CTA_ROWS rays per CTA (one 16 x 2 strip per warp) */
#endif
#ifndef TRACE_FOLD_CAM
#define TRACE_FOLD_CAM 1                /* 1: the horizon sample's camera products come from the kernel parameters (TraceK::camf) */
#endif
#ifndef TRACE_FOLD_T
#define TRACE_FOLD_T 1                  /* 1: t to the cell's far planes as ONE FFMA per axis: ((cell + fo) * 2^m) * (1 / (res * dir)) + (c / dir) */
#endif
#ifndef TRACE_FOLD_Z
#define TRACE_FOLD_Z 1                  /* 1: t to the depth plane as ONE FFMA, sz * (1 / dz) + (-oz / dz); for dz <= 0 the pair is (0, MAX_T) */
#endif
#ifndef TRACE_PINNED_FLAGLESS
#define TRACE_PINNED_FLAGLESS 0         /* 1: the pinned loop ends on one condition (no exit flags held in predicates): 59 -> 56 instructions per step,
                                           0.3317 -> 0.3280 ms at 4K with identical pass-level parity at 1280x720 (profiles/r02ak_...).  NOT the default: the
                                           full GPU suite was not rerun with it before the round's GPU budget ran out */
#endif
#ifndef TRACE_SS_PARAM
#define TRACE_SS_PARAM 1                /* 1: the level-0 extent as floats in the kernel parameters (no per-step I2FP re-materialisation) */
#endif
namespace vkr {

#define TR_MAGIC0 12582912.0f          /* 1.5 * 2^23 */
#define TR_MAGIC0_BITS 0x4B400000

struct RayK {                          // per-ray constants of the march
  float ox, oy, oz, dx, dy, dz;        // origin, direction (uv, depth)
  float idx, idy, idz;                 // 1 / direction (MAX_T for 0)
  float fox, foy;                      // floor offsets (0 or 1)
  float cx, cy;                        // uv_offset - origin
  float t_exit;                        // the fast loops are valid while t < t_exit
#if TRACE_FOLD_Z
  float ozi;                           // -oz / dz (MAX_T with idz = 0 when dz <= 0: the depth plane is never the nearest)
#endif
#if TRACE_FOLD_T
  float kx, ky, cxi, cyi;              // 1 / (res0 * direction), (uv_offset - origin) / direction; during the pinned loop cxi, cyi include fo * k
#endif
};

// Horizon sample (trace.comp:258-262): v = reconstruct(uv, sz) - view_vec at the ray's CURRENT uv, h = max(h, n.v / |v|) if
// |v| < 0.3.  |v| >= |v.z|, so a sample whose view z is 0.3 or more away is dropped before the rest is evaluated; a NaN sz
// ("no sample owed") drops out at the same test.
__device__ __forceinline__ void horizon_sample(const f3& vv, const f3& n, const CamF& camf, float m2at, float m2tg, float px, float py, float sz, float& h) {
  const float z = linearize_fast(sz, camf);
  const float vz = z - vv.z;
  if (fabsf(vz) < 0.3f) {
    const float vx = fmaf(fmaf(m2at, px, camf.aspect_tg), z, -vv.x), vy = fmaf(fmaf(m2tg, py, camf.tg), z, -vv.y);
    const float d2 = fmaf(vz, vz, fmaf(vy, vy, vx * vx));
    if (d2 < 0.09f) h = fmaxf(h, fmaf(n.z, vz, fmaf(n.y, vy, n.x * vx)) * rsqrt_fast(d2));
  }
}

// one step on level m (magic = 1.5 * 2^(23+m)); returns above_surface
template <bool PINNED>
__device__ __forceinline__ bool fast_step(const RayK& r, const uint4* __restrict__ lvl, float ssx, float ssy, float rinvx0, float rinvy0, float magic,
                                          float px, float py, float& t, bool& skipped, float& sz_out) {
  const float mx = __fmaf_rd(ssx, px, magic), my = __fmaf_rd(ssy, py, magic);
  const int mb = __float_as_int(magic);
  const int xi = __float_as_int(mx) - mb, yi = __float_as_int(my) - mb;            // cell index on this level (>= 0 inside the screen)
  const uint4 L = PINNED ? lvl[0] : lvl[(mb >> 23) - 150];
  const uint32_t* base = (const uint32_t*)(((unsigned long long)L.y << 32) | L.x);
  float sz = 0.0f;
  if (PINNED || ((unsigned)xi < (L.w & 0xFFFFu) && (unsigned)yi < (L.w >> 16))) {  // level 0: res is the integer extent, inside the screen = inside the level
    const uint32_t texel = (uint32_t)yi * L.z + (uint32_t)xi;                        // 32-bit index, widened once
    sz = d24_to_float(__ldg(base + texel));
  }
#if TRACE_FOLD_T
  // level 0: cell * k + (fo * k + c / dir), the bracket held in cxi / cyi.  Level m: (cell + fo) * 2^m = mx - (magic - fo * 2^m),
  // and magic - 2^m is magic's bit pattern minus one: exact
  const float qx = PINNED ? mx - TR_MAGIC0 : mx - __int_as_float(mb - (r.dx < 0 ? 0 : 1)), qy = PINNED ? my - TR_MAGIC0 : my - __int_as_float(mb - (r.dy < 0 ? 0 : 1));
  const float tx = fmaf(qx, r.kx, r.cxi), ty = fmaf(qy, r.ky, r.cyi);
#else
  const float p2 = PINNED ? 1.0f : __int_as_float(mb - (TR_MAGIC0_BITS - 0x3F800000));   // 2^m
  const float ex = fmaf(fmaf(p2, r.fox, mx - magic), rinvx0, r.cx), ey = fmaf(fmaf(p2, r.foy, my - magic), rinvy0, r.cy);   // xy_plane - origin
  const float tx = ex * r.idx, ty = ey * r.idy;
#endif
#if TRACE_FOLD_Z
  const float tz = fmaf(sz, r.idz, r.ozi);
#else
  float tz = (sz - r.oz) * r.idz;
  tz = r.dz > 0 ? tz : MAX_T_FLOAT;
#endif
  const float t_min = fminf(fminf(tx, ty), tz);
  const bool above = sz > fmaf(t, r.dz, r.oz);
  skipped = above && (t_min != tz);
  t = above ? t_min : t;
  sz_out = sz;
  return above;
}

template <bool LAYERED>
__global__ void __launch_bounds__(16 * TRACE_CTA_ROWS, TRACE_MIN_CTAS) sssr_trace2_kernel(TraceK k0, const TraceK* __restrict__ kv, LaunchX lx, PyramidView P, Mip normal, Mip material,
                                                          const float4* __restrict__ halton, Mip pdf, Mip out_rays, Mip out_occl, const float* __restrict__ srgb, TraceDims dims) {
  __shared__ uint4 lvl[16];
  TraceK kl;
  unsigned long long poff = 0;
  if (LAYERED) {
    const int z = blockIdx.z;
    kl = kv[z];
    poff = (unsigned long long)z * P.mip[0].ls;
    normal = layer_of(normal, z); material = layer_of(material, z); pdf = layer_of(pdf, z); out_rays = layer_of(out_rays, z); out_occl = layer_of(out_occl, z);
  }
  const TraceK& k = LAYERED ? kl : k0;
  if (threadIdx.x < 16) {
    const int m = threadIdx.x;
    uint4 e = make_uint4(0u, 0u, 0u, 0u);                   // levels past the view: extent 0 -> texelFetch returns 0
    if (m < P.count) {
      const unsigned long long b = (unsigned long long)(P.mip[m].p + poff);
      e = make_uint4((uint32_t)b, (uint32_t)(b >> 32), P.mip[m].pitch >> 2, (uint32_t)P.mip[m].w | ((uint32_t)P.mip[m].h << 16));
    }
    lvl[m] = e;
  }
  __syncthreads();
  const int px = blockIdx.x * 16 + (threadIdx.x & 15), py = VKR_BLOCK_Y * TRACE_CTA_ROWS + (threadIdx.x >> 4);
  if (px >= out_rays.w || py >= out_rays.h) return;
  const float tsx = (float)out_rays.w, tsy = (float)out_rays.h;
  const float su = ((float)px + 0.5f) / tsx, sv = ((float)py + 0.5f) / tsy;                  // :44
  const CamK cam = k.cam;
#if TRACE_FOLD_CAM
  const CamF camf = k.camf;
  const float m2at = k.m2at, m2tg = k.m2tg;
#else
  const CamF camf = make_camf(cam);
  const float m2at = -2.0f * camf.aspect_tg, m2tg = -2.0f * camf.tg;
#endif

  float roughness = sample_srgb8(material, su, sv, srgb).y;                                     // :49-53
  float mg = gmix(0.0f, k.max_roughness, roughness);
  roughness = FM(mg, mg);

  Mip d0 = P.mip[0];
  d0.p += poff;
  const float pixel_depth = sample_d24(d0, su, sv);
  const uint32_t base_index = (uint32_t)(rand_hash(su, sv) * 128.0f);                           // :61-63
  const uint32_t index = (base_index + k.frame_random) & 127u;
  const float4 hv = __ldg(halton + index);
  f3 pixel_normal, R, origin, direction;
  const f3 view_vec = reconstruct_view_vec(su, sv, pixel_depth, cam);                           // exact: one side of the oracle-bit horizon difference
#if TRACE_FAST_SETUP
  {
    pixel_normal = norm3f(mul_m3f(k.normal_mat, sample_gbuffer_normal_fast(normal, su, sv)));
    f3 tangent = get_tangent(pixel_normal);
    f3 bitangent = norm3f(cross3(pixel_normal, tangent));
    tangent = norm3f(cross3(bitangent, pixel_normal));
    const f3 w0 = -norm3f(view_vec);
    f3 view_dir = mk3(dot3f(w0, tangent), dot3f(w0, bitangent), dot3f(w0, pixel_normal));
    f3 brdf_norm = sampleGGXVNDF_fast(view_dir, roughness, roughness, hv.x, hv.y);
    f3 N = mk3(fmaf(brdf_norm.x, tangent.x, fmaf(brdf_norm.y, bitangent.x, brdf_norm.z * pixel_normal.x)),
               fmaf(brdf_norm.x, tangent.y, fmaf(brdf_norm.y, bitangent.y, brdf_norm.z * pixel_normal.y)),
               fmaf(brdf_norm.x, tangent.z, fmaf(brdf_norm.y, bitangent.z, brdf_norm.z * pixel_normal.z)));
    const float two_ndi = 2.0f * dot3f(N, view_vec);
    R = mk3(fmaf(-two_ndi, N.x, view_vec.x), fmaf(-two_ndi, N.y, view_vec.y), fmaf(-two_ndi, N.z, view_vec.z));
    origin = project_fast(mk3(fmaf(0.001f, pixel_normal.x, view_vec.x), fmaf(0.001f, pixel_normal.y, view_vec.y), fmaf(0.001f, pixel_normal.z, view_vec.z)), cam);
    origin.z -= 0.0001f;
    f3 e = project_fast(mk3(view_vec.x + R.x, view_vec.y + R.y, view_vec.z + R.z), cam);
    direction = mk3(e.x - origin.x, e.y - origin.y, e.z - origin.z);
    const float s = (1.0f - origin.z) * rcp_fast(direction.z);
    direction = mk3(direction.x * s, direction.y * s, direction.z * s);
  }
#else
  {
    pixel_normal = norm3(mul_m4(k.normal_mat, sample_gbuffer_normal(normal, su, sv), 0.f));
    f3 tangent = get_tangent(pixel_normal);
    f3 bitangent = norm3(cross3(pixel_normal, tangent));
    tangent = norm3(cross3(bitangent, pixel_normal));
    const f3 w0 = -norm3(view_vec);
    f3 view_dir = mk3(dot3(w0, tangent), dot3(w0, bitangent), dot3(w0, pixel_normal));
    f3 brdf_norm = sampleGGXVNDF(view_dir, roughness, roughness, hv.x, hv.y);
    f3 N = brdf_norm.x * tangent + brdf_norm.y * bitangent + brdf_norm.z * pixel_normal;
    R = reflect3(view_vec, N);
    origin = project_view_vec(view_vec + 0.001f * pixel_normal, cam);                            // :79-84
    origin.z = FS(origin.z, 0.0001f);
    direction = project_view_vec(view_vec + R, cam) - origin;
    direction = direction * (FS(1.f, origin.z) / direction.z);
  }
#endif
  // ---- hierarchical_raymarch_find_hor, trace.comp:206-268 ------------------------------------
#if TRACE_SS_PARAM
  const float ssx = dims.ssx, ssy = dims.ssy;
#else
  const float ssx = (float)d0.w, ssy = (float)d0.h;
#endif
  const float rinvx0 = 1.0f / ssx, rinvy0 = 1.0f / ssy;
  RayK r;
  r.ox = origin.x; r.oy = origin.y; r.oz = origin.z; r.dx = direction.x; r.dy = direction.y; r.dz = direction.z;
#if TRACE_FAST_SETUP
  r.idx = direction.x != 0 ? rcp_fast(direction.x) : MAX_T_FLOAT; r.idy = direction.y != 0 ? rcp_fast(direction.y) : MAX_T_FLOAT;
  r.idz = direction.z != 0 ? rcp_fast(direction.z) : MAX_T_FLOAT;
#else
  r.idx = direction.x != 0 ? 1.0f / direction.x : MAX_T_FLOAT; r.idy = direction.y != 0 ? 1.0f / direction.y : MAX_T_FLOAT;
  r.idz = direction.z != 0 ? 1.0f / direction.z : MAX_T_FLOAT;
#endif
  const float uox = direction.x < 0 ? -(0.005f / ssx) : (0.005f / ssx), uoy = direction.y < 0 ? -(0.005f / ssy) : (0.005f / ssy);
  r.fox = direction.x < 0 ? 0.f : 1.f; r.foy = direction.y < 0 ? 0.f : 1.f;
  r.cx = uox - origin.x; r.cy = uoy - origin.y;
#if TRACE_FOLD_Z
  if (direction.z > 0) r.ozi = -origin.z * r.idz; else { r.idz = 0.0f; r.ozi = MAX_T_FLOAT; }
#endif
#if TRACE_FOLD_T
  r.kx = rinvx0 * r.idx; r.ky = rinvy0 * r.idy;
  r.cxi = fmaf(r.fox, r.kx, r.cx * r.idx); r.cyi = fmaf(r.foy, r.ky, r.cy * r.idy);       // pinned-loop form
#endif
  {                                                    // the ray stays inside [2^-20, 1 - 2^-20]^2 while t < t_exit
    const float eps = 9.5367431640625e-07f;
    const float bx = direction.x < 0 ? eps : 1.0f - eps, by = direction.y < 0 ? eps : 1.0f - eps;
    const float ex = direction.x != 0 ? (bx - origin.x) * r.idx : MAX_T_FLOAT, ey = direction.y != 0 ? (by - origin.y) * r.idy : MAX_T_FLOAT;
    r.t_exit = fminf(ex, ey);
    const bool inside = origin.x > eps && origin.x < 1.0f - eps && origin.y > eps && origin.y < 1.0f - eps;
    if (!inside) r.t_exit = -1.0f;
  }
  float t;
  {                                                    // initial_advance_ray, screen_trace.glsl:8-15
    const float ex = fmaf(floorf(ssx * origin.x) + r.fox, rinvx0, r.cx), ey = fmaf(floorf(ssy * origin.y) + r.foy, rinvy0, r.cy);
    t = fminf(ex * r.idx, ey * r.idy);
  }
  // The horizon sample of a step is taken at the position the ray has AFTER that step (:256-262).  It is evaluated at the
  // top of the next step, where that position's uv is needed anyway; `owed` carries the step's depth (NaN = nothing owed).
  const float NOTHING = __int_as_float(0x7FC00000);
  float h = 0.f, owed = NOTHING;
  int i = 0;
  int mip = 0;
  // Steps 0..14 stay on level 0 (:250).  A ray found below the surface there never moves again (t frozen, the same
  // horizon sample every step) and leaves the march at step 15: it is finished.
#if TRACE_PINNED_FLAGLESS
  bool above0 = true;                                  // one loop condition, no exit flags to keep in predicates
  while (above0 && i < 15 && t < r.t_exit) {
    const float cpx = fmaf(t, r.dx, r.ox), cpy = fmaf(t, r.dy, r.oy);
    horizon_sample(view_vec, pixel_normal, camf, m2at, m2tg, cpx, cpy, owed, h);
    bool skipped;
    above0 = fast_step<true>(r, lvl, ssx, ssy, rinvx0, rinvy0, TR_MAGIC0, cpx, cpy, t, skipped, owed);
    ++i;
  }
  bool done = !above0;                                 // (i counts the finishing step too; it is not read again for a finished ray)
#else
  bool done = false;
  for (; i < 15 && t < r.t_exit; ++i) {
    const float cpx = fmaf(t, r.dx, r.ox), cpy = fmaf(t, r.dy, r.oy);
    horizon_sample(view_vec, pixel_normal, camf, m2at, m2tg, cpx, cpy, owed, h);
    bool skipped;
    const bool above = fast_step<true>(r, lvl, ssx, ssy, rinvx0, rinvy0, TR_MAGIC0, cpx, cpy, t, skipped, owed);
    if (!above) { done = true; break; }
  }
#endif
#if TRACE_FOLD_T
  r.cxi = fmaf(r.dx < 0 ? 0.f : -1.f, r.kx, r.cxi); r.cyi = fmaf(r.dy < 0 ? 0.f : -1.f, r.ky, r.cyi);   // back to c / dir for the levels above 0
#endif
  if (!done && i >= 15) {
    float magic = TR_MAGIC0;
    while (i < 80 && magic >= TR_MAGIC0 && t < r.t_exit) {
      const float cpx = fmaf(t, r.dx, r.ox), cpy = fmaf(t, r.dy, r.oy);
      if (owed == owed) horizon_sample(view_vec, pixel_normal, camf, m2at, m2tg, cpx, cpy, owed, h);
      bool skipped; float sz;
      fast_step<false>(r, lvl, ssx, ssy, rinvx0, rinvy0, magic, cpx, cpy, t, skipped, sz);
      magic = __int_as_float(__float_as_int(magic) + (skipped ? 0x00800000 : -0x00800000));
      ++i;
      owed = magic <= 2.0f * TR_MAGIC0 ? sz : NOTHING;
    }
    mip = (__float_as_int(magic) >> 23) - 150;
  }
  // Slow tail: the ray is at the screen border or outside.  Literal texelFetch semantics: index = trunc(res * uv) (so the
  // strip (-1, 0) still reads cell 0), out of range = 0; floor() for the plane (screen_trace.glsl:19).
  while (!done && i < 80 && mip >= 0) {
    const float S = __int_as_float((127 + mip) << 23), s = __int_as_float((127 - mip) << 23);   // 2^mip, 2^-mip
    const float pxx = fmaf(t, r.dx, r.ox), pyy = fmaf(t, r.dy, r.oy);
    if (owed == owed) horizon_sample(view_vec, pixel_normal, camf, m2at, m2tg, pxx, pyy, owed, h);
    const float mpx = (ssx * s) * pxx, mpy = (ssy * s) * pyy;
    const int xi = (int)mpx, yi = (int)mpy;
    const uint4 L = lvl[min(mip, 15)];
    const bool oobx = (unsigned)xi >= (L.w & 0xFFFFu), ooby = (unsigned)yi >= (L.w >> 16);
    if ((oobx && (pxx >= 1.0f || mpx <= -1.0f)) || (ooby && (pyy >= 1.0f || mpy <= -1.0f))) {
      // Out of range along one axis on this level AND on every finer one (uv >= 1 with cell >= extent here: the cell index
      // at least doubles per level while the extent at most doubles; res * uv <= -1: ever more negative as res doubles).
      // Every remaining fetch returns 0, which is never above the ray: t stays, the level counts down to -1, and the same
      // sample (depth 0) is offered to the horizon at each level <= 1 -- once is enough, if the step budget lets the march
      // get that far: the first such step is number max(mip - 1, 1) from here (step 1 at once while the level is pinned).
      const int first = (i < 15) ? 1 : max(mip - 1, 1);
      owed = (i + first <= 80) ? 0.0f : NOTHING;
      done = true;
      break;
    }
    float sz = 0.0f;
    if (!oobx && !ooby)
      sz = d24_to_float(__ldg((const uint32_t*)(((unsigned long long)L.y << 32) | L.x) + ((uint32_t)yi * L.z + (uint32_t)xi)));
#if TRACE_FOLD_T
    const float tx = fmaf((floorf(mpx) + (r.dx < 0 ? 0.f : 1.f)) * S, r.kx, r.cxi), ty = fmaf((floorf(mpy) + (r.dy < 0 ? 0.f : 1.f)) * S, r.ky, r.cyi);
#else
    const float ex = fmaf((floorf(mpx) + r.fox) * S, rinvx0, r.cx), ey = fmaf((floorf(mpy) + r.foy) * S, rinvy0, r.cy);
    const float tx = ex * r.idx, ty = ey * r.idy;
#endif
#if TRACE_FOLD_Z
    const float tz = fmaf(sz, r.idz, r.ozi);
#else
    float tz = (sz - r.oz) * r.idz;
    tz = r.dz > 0 ? tz : MAX_T_FLOAT;
#endif
    const float t_min = fminf(fminf(tx, ty), tz);
    const bool above = sz > fmaf(t, r.dz, r.oz);
    const bool skipped = above && (t_min != tz);
    t = above ? t_min : t;
    if (i >= 15) mip += skipped ? 1 : -1;
    ++i;
    owed = mip <= 1 ? sz : NOTHING;
  }
  const f3 out_ray = mk3(fmaf(t, r.dx, r.ox), fmaf(t, r.dy, r.oy), fmaf(t, r.dz, r.oz));
  if (owed == owed) horizon_sample(view_vec, pixel_normal, camf, m2at, m2tg, out_ray.x, out_ray.y, owed, h);
  bool valid_hit = true;                              // (i <= max) is always true, App. A.6 #8

  {                                                                                              // :94-99
    float sx = fabsf(out_ray.x - origin.x) * tsx, sy = fabsf(out_ray.y - origin.y) * tsy;
    if (fmaxf(sx, sy) < 2.0f) valid_hit = false;
  }
  if (valid_hit) {                                                                               // :101-109
#if TRACE_FAST_SETUP
    f3 hit_normal = mul_m3f(k.normal_mat, sample_gbuffer_normal_fast(normal, out_ray.x, out_ray.y));
    if (dot3f(hit_normal, R) > 0 || dot3f(pixel_normal, R) < 0) valid_hit = false;
#else
    f3 hit_normal = mul_m4(k.normal_mat, sample_gbuffer_normal(normal, out_ray.x, out_ray.y), 0.f);
    if (dot3(hit_normal, R) > 0 || dot3(pixel_normal, R) < 0) valid_hit = false;
#endif
  }
  if (valid_hit) {                                                                               // :111-118
    float hit_depth = sample_d24(d0, out_ray.x, out_ray.y);
    float hit_z = linearize_fast(hit_depth, camf);
    float ray_z = linearize_fast(out_ray.z, camf);
    if (ray_z > hit_z + 0.3f || ray_z < hit_z - 0.1f) valid_hit = false;
  }
  {                                                                                              // :120-121, RGBA16_UNORM
    uint32_t lo = to_unorm(out_ray.x, 65535.f) | (to_unorm(out_ray.y, 65535.f) << 16);
    uint32_t hi = to_unorm(out_ray.z, 65535.f) | (to_unorm(valid_hit ? pixel_depth : 1.0f, 65535.f) << 16);
    *((uint2*)(out_rays.p + (size_t)py * out_rays.pitch) + px) = make_uint2(lo, hi);
  }
  {                                                                                              // :123-139, value-only
#if TRACE_FAST_SETUP
    const f3 w0 = -norm3f(view_vec);                  // recomputed: not kept alive across the march
    f3 slice_normal = norm3f(cross3(w0, R));
    const float pn = dot3f(pixel_normal, slice_normal);
    f3 normal_projected = mk3(fmaf(-pn, slice_normal.x, pixel_normal.x), fmaf(-pn, slice_normal.y, pixel_normal.y), fmaf(-pn, slice_normal.z, pixel_normal.z));
    f3 X = norm3f(cross3(slice_normal, w0));
    const float np2 = dot3f(normal_projected, normal_projected);
    const float rnp = rsqrt_fast(np2);
    float n = VKR_PI / 2.0f - acosf(dot3f(normal_projected, X) * rnp);
    float hh = acosf(h);
    hh = gmin(n + gmin(hh - n, VKR_PI / 2.0f), hh);
    float pdfv = sampleGGXdirPDF(pdf, w0, pixel_normal, R, roughness);
    float occlusion = ((1.0f / VKR_PI) * 0.25f) * (np2 * rnp) * gmax(-__cosf(2.f * hh - n) + __cosf(n) + 2.f * hh * __sinf(n), 0.f);
#else
    const f3 w0 = -norm3(view_vec);
    f3 slice_normal = norm3(cross3(w0, R));
    f3 normal_projected = pixel_normal - dot3(pixel_normal, slice_normal) * slice_normal;
    f3 X = norm3(cross3(slice_normal, w0));
    float n = FS(VKR_PI / 2.0f, acosf(dot3(norm3(normal_projected), X)));
    float hh = acosf(h);
    hh = gmin(FA(n, gmin(FS(hh, n), VKR_PI / 2.0f)), hh);
    float pdfv = sampleGGXdirPDF(pdf, w0, pixel_normal, R, roughness);
    float occlusion = FM(FM(FM(1.0f / VKR_PI, len3(normal_projected)), 0.25f), gmax(FA(FA(-cosf(FS(FM(2.f, hh), n)), cosf(n)), FM(FM(2.f, hh), sinf(n))), 0.f));
#endif
    float result = isnan(occlusion) ? 0.0f : occlusion;
#ifdef VKR_TRACE_STATS
    *((uint2*)(out_occl.p + (size_t)py * out_occl.pitch) + px) = make_uint2(pack_half2(result, pdfv), pack_half2((float)i, (float)(done ? -1 : mip)));
#else
    *((uint2*)(out_occl.p + (size_t)py * out_occl.pitch) + px) = make_uint2(pack_half2(result, pdfv), 0u);
#endif
  }
}

}  // namespace vkr
using namespace vkr;

static TraceK make_trace_k(const VkrTraceParams* p) {
  TraceK k;
  k.normal_mat = make_mat(p->normal_mat);
  k.cam = make_cam(p->fovy, p->aspect, p->znear, p->zfar);
  k.frame_random = p->frame_random; k.max_roughness = p->max_roughness;
  k.camf.aspect_tg = k.cam.aspect * k.cam.tg; k.camf.tg = k.cam.tg;      // make_camf's roundings: one multiply / subtract each
  k.camf.nf = k.cam.znear * k.cam.zfar; k.camf.fmn = k.cam.zfar - k.cam.znear; k.camf.zfar = k.cam.zfar;
  k.m2at = -2.0f * k.camf.aspect_tg; k.m2tg = -2.0f * k.cam.tg;
  return k;
}

extern "C" int vkr_sssr_trace_ex(const VkrTraceParams* p, const VkrImage* depth, const VkrImage* dnormal, const VkrImage* material,
                                 const float* halton_vec4x128, const VkrImage* pdf_lut, const VkrImage* out_rays,
                                 const VkrImage* out_occlusion, const VkrLaunchExt* ext, void* stream) {
  if (!p || !halton_vec4x128 || !image_ok(depth, VKR_FORMAT_D24_UNORM_S8_UINT) || depth->mip_count < 2 ||
      !image_ok(dnormal, VKR_FORMAT_R16G16_UNORM) || !image_ok(material, VKR_FORMAT_R8G8B8A8_SRGB) || !image_ok(pdf_lut, VKR_FORMAT_R32_SFLOAT) ||
      !image_ok(out_rays, VKR_FORMAT_R16G16B16A16_UNORM) || !image_ok(out_occlusion, VKR_FORMAT_R16G16B16A16_SFLOAT)) {
    set_error("sssr_trace: null argument or unexpected format");
    return VKR_ERR_INVALID_ARGUMENT;
  }
  if (out_rays->width != out_occlusion->width || out_rays->height != out_occlusion->height) {
    set_error("sssr_trace: rays and occlusion sizes differ");
    return VKR_ERR_SIZE_MISMATCH;
  }
  ExtInfo x;
#if TRACE_V2
  int rc = resolve_ext(ext, out_rays->height, TRACE_CTA_ROWS, &x, "sssr_trace");
#else
  int rc = resolve_ext(ext, out_rays->height, 8, &x, "sssr_trace");
#endif
  if (rc) return rc;
  cudaStream_t st = (cudaStream_t)stream;
  TraceK k = make_trace_k(p);
  const TraceK* kv = nullptr;
  if (x.layers > 1) {
    if (!x.prepared()) {
      std::vector<TraceK> ks(x.layers);
      for (int i = 0; i < x.layers; i++) ks[i] = make_trace_k(p + i);
      if ((rc = upload_params(ks.data(), sizeof(TraceK), x.layers, x.scratch(), st, "sssr_trace"))) return rc;
    }
    kv = (const TraceK*)x.scratch() + x.first();
    if (x.prepare_only()) return VKR_OK;
  }
  PyramidView P;
  P.count = (int)depth->mip_count - 1;
  for (int i = 0; i < VKR_MAX_MIPS; i++) P.mip[i] = (i < P.count) ? make_mip(depth, i + 1, x, 0) : Mip{nullptr, 0, 0, 0, 0};
  const float* srgb = srgb_decode_table_device();
  if (!srgb) { set_error("sssr_trace: sRGB table upload failed"); return VKR_ERR_CUDA; }
  dim3 grid((out_rays->width + 15) / 16, x.yblocks, x.layers);
#if TRACE_V2
  auto kern = kv ? sssr_trace2_kernel<true> : sssr_trace2_kernel<false>;
#else
  auto kern = kv ? sssr_trace_kernel<true> : sssr_trace_kernel<false>;
#endif
  const int threads = TRACE_V2 ? 16 * TRACE_CTA_ROWS : 128;
  kern<<<grid, threads, 0, st>>>(k, kv, LaunchX{x.yb0}, P, make_mip(dnormal, 0, x, 1), make_mip(material, 0, x, 2), (const float4*)halton_vec4x128,
                             make_mip(pdf_lut, 0, x, 3), make_mip(out_rays, 0, x, 4), make_mip(out_occlusion, 0, x, 5), srgb,
                             TraceDims{(float)P.mip[0].w, (float)P.mip[0].h});
  return check_launch("sssr_trace_kernel");
}

extern "C" int vkr_sssr_trace(const VkrTraceParams* p, const VkrImage* depth, const VkrImage* dnormal, const VkrImage* material,
                              const float* halton_vec4x128, const VkrImage* pdf_lut, const VkrImage* out_rays,
                              const VkrImage* out_occlusion, void* stream) {
  return vkr_sssr_trace_ex(p, depth, dnormal, material, halton_vec4x128, pdf_lut, out_rays, out_occlusion, nullptr, stream);
}
