// ssr_blur.cu — roughness-sized bilateral Gaussian blur of the resolved reflections plus the
// temporal blend.  Replaces task "SSSR_blur" (reference src/advanced_ssr.cpp:371-438,
// shaders/advanced_ssr/blur.comp:31-118).
//
// The reference shader fetches, for each of up to (2*11+1)^2 = 529 taps per pixel, one depth
// texel, one bilinear normal (4 texels + octahedral decode + normalize) and one colour texel.
// All three depend only on the tap POSITION, so a CTA of 32x16 pixels stages them once per
// position for its tile + halo (halo = the tile's largest radius) in shared memory.
// The bilateral weights are per tap, so the filter is not separable.
//
// The tap loop is what the pass costs, and it is bound by issue slots, so it runs on packed fp32 (f32x2.cuh):
//   * the staged data is laid out in ROW PAIRS — {nx, nx' | ny, ny'}, {nz, nz' | depth, depth'}, {r, r' | g, g'}, {b, b'}
//     for rows 2P and 2P+1 of the staged window — so one LDS.128 delivers two taps as ready-made register pairs;
//   * every thread owns TWO vertically adjacent pixels and walks their common column once (rows -r .. r+1, rounded out
//     to whole pairs): 4 loads and 2 x (10 packed + 4 scalar) FP instructions weight 2 taps for 2 pixels —
//     8 issue slots per pixel and tap against 17 for the scalar one-pixel-per-thread form;
//   * the Gaussian along the column is a recurrence over row pairs: G(j+2) = G(j) q(j), q(j+2) = q(j) c.
// Warps iterate the window of their largest radius in lockstep (pixels with a smaller radius carry a zero Gaussian
// outside their own window), which keeps the shared-memory reads conflict-free.  Each accumulator is two partial sums
// (even and odd rows of the staged window) added at the end: value-only arithmetic, see the precision policy.
#include <vector>
#include "common.cuh"
#include "f32x2.cuh"

namespace vkr {

struct BlurK {
  Mat4 inverse_camera, prev_inverse_camera;
  CamK cam;
  float max_roughness;
  uint32_t accumulate, disable_blur;
};

#ifndef BLUR_MIN_CTAS
#define BLUR_MIN_CTAS 3
#endif
#define BLUR_TW 32
#define BLUR_TH 16
#define BLUR_RMAX 11
#define BLUR_SPAN_X (BLUR_TW + 2 * BLUR_RMAX)
#define BLUR_PAIRS (BLUR_TH / 2 + BLUR_RMAX)
#define BLUR_PLANE (BLUR_SPAN_X * BLUR_PAIRS)
// BLUR_BLUE_HALF: the {b, b'} plane as half2 (the colours are unorm8 texels: the fp16 error is 0.06 LSB).  57 456 -> 53 352
// bytes per CTA, which is what lets a FOURTH CTA fit an SM's 228 KB (4 x (56 KB + 1 KB reserved)); with the 64-register cap
// of BLUR_MIN_CTAS = 4 the barrier-separated phases of the CTAs cover each other's load latency better.
#ifndef BLUR_BLUE_HALF
#define BLUR_BLUE_HALF 0
#endif
#define BLUR_SMEM_BYTES (BLUR_PLANE * (3 * 16 + (BLUR_BLUE_HALF ? 4 : 8)))

// blur.comp:110-115 — exact depth footprint and z denominator (thresholded downstream), value-only otherwise
__device__ __forceinline__ f3 world_pos_lod(const Mip& depth1, const Mat4& inv, float u, float v, const CamF& cam) {
  return mul_m4f(inv, reconstruct_fast(u, v, sample_d24(depth1, u, v), cam));
}

// Per-pixel state of the tap loop; every f2 holds the values for the even and the odd row of the current row pair.
struct BlurPx {
  f2 ncd;            // -centre depth, duplicated
  float bk;          // 1000 / centre depth
  f2 nx, ny, nz;     // centre normal, duplicated
  f2 g, q, c8;       // Gaussian of the two rows, G(j+2)/G(j), and that ratio's own (constant) ratio
  f2 cr, cg, cb, ws;
};
struct BlurTap { ulonglong2 A, B, C; unsigned long long D; };

// bilateral x normal weight of the two taps (blur.comp:66-69): max(1 - |dz| 1000/z, 0) * max(n . n', 0)
__device__ __forceinline__ f2 blur_weight(const BlurPx& s, const BlurTap& t) {
  float dl, dh, nl, nh;
  unpk(add2(f2{t.B.y}, s.ncd), dl, dh);
  const float bl = __saturatef(fmaf(-fabsf(dl), s.bk, 1.0f)), bh = __saturatef(fmaf(-fabsf(dh), s.bk, 1.0f));
  unpk(fma2(s.nz, f2{t.B.x}, fma2(s.ny, f2{t.A.y}, mul2(s.nx, f2{t.A.x}))), nl, nh);
  return pk(__saturatef(bl * nl), __saturatef(bh * nh));       // b in [0,1], n <= 1: clamping the product clamps n at 0
}
__device__ __forceinline__ void blur_accumulate(BlurPx& s, const BlurTap& t, f2 w) {
  s.cr = fma2(f2{t.C.x}, w, s.cr); s.cg = fma2(f2{t.C.y}, w, s.cg); s.cb = fma2(f2{t.D}, w, s.cb);
  s.ws = add2(s.ws, w);
}

template <bool SAME, bool LAYERED>
__global__ void __launch_bounds__(256, BLUR_MIN_CTAS) sssr_blur_kernel(BlurK k0, const BlurK* __restrict__ kv, LaunchX lx, Dim dim, Mip depth1, Mip normal_tex, Mip refl,
                                                           Mip material_tex, Mip history, Mip velocity, Mip prev_depth1, Mip out, const float* __restrict__ srgb, PeerOut po) {
  extern __shared__ float4 s_blur[];
  float4* const sA = s_blur;                          // {nx, nx' | ny, ny'}
  float4* const sB = sA + BLUR_PLANE;                 // {nz, nz' | depth, depth'}
  float4* const sC = sB + BLUR_PLANE;                 // {r, r' | g, g'}
#if BLUR_BLUE_HALF
  __half2* const sD = (__half2*)(sC + BLUR_PLANE);    // {b, b'} as half2
#else
  float2* const sD = (float2*)(sC + BLUR_PLANE);      // {b, b'}
#endif
  __shared__ int s_rmax;
  BlurK kl;
  if (LAYERED) {
    const int z = blockIdx.z;
    kl = kv[z];
    depth1 = layer_of(depth1, z); normal_tex = layer_of(normal_tex, z); refl = layer_of(refl, z); material_tex = layer_of(material_tex, z);
    history = layer_of(history, z); velocity = layer_of(velocity, z); prev_depth1 = layer_of(prev_depth1, z); out = layer_of(out, z);
  }
  const BlurK& k = LAYERED ? kl : k0;

  // warp w owns rows 2w and 2w+1 of the tile, lane = column
  const int tx = threadIdx.x & 31, ty = (threadIdx.x >> 5) * 2;
  const int px = blockIdx.x * BLUR_TW + tx, py0 = VKR_BLOCK_Y * BLUR_TH + ty;
  const float su = pixel_uv(px, dim.w, dim.rw);
  BlurPx s[2];
  float sigma[2];
  int r[2];
  bool inside[2];

  // per-pixel radius (blur.comp:39-53).  floor() makes this the one discontinuous quantity of the pass: unfused.
#pragma unroll
  for (int p = 0; p < 2; p++) {
    inside[p] = px < out.w && py0 + p < out.h;
    sigma[p] = 0.35f;
    r[p] = 0;
    if (inside[p]) {
      float roughness = sample_srgb8(material_tex, su, pixel_uv(py0 + p, dim.h, dim.rh), srgb).y;
      roughness = gmix(0.0f, k.max_roughness, roughness);
      sigma[p] = gmix(0.4f, 4.f, roughness);
      if (k.disable_blur != 0) sigma[p] = 0.35f;
      r[p] = min(max((int)floorf(FS(FM(3.f, sigma[p]), 0.01f)), 0), BLUR_RMAX);
    }
  }
  if (threadIdx.x == 0) s_rmax = 0;
  __syncthreads();
  const int wr = __reduce_max_sync(0xFFFFFFFFu, max(r[0], r[1]));
  if ((threadIdx.x & 31) == 0) atomicMax(&s_rmax, wr);
  __syncthreads();
  const int R = s_rmax;
  const int span_x = BLUR_TW + 2 * R, pairs = BLUR_TH / 2 + R;

  // stage {normal, depth} and colour for every tap position of the tile, two rows per item
  const int ox = blockIdx.x * BLUR_TW - R, oy = VKR_BLOCK_Y * BLUR_TH - R;
  const uint32_t span_magic = div_magic((uint32_t)span_x);             // span_x * pairs <= 54 * 19
  for (int t = threadIdx.x; t < span_x * pairs; t += 256) {
    const int P = div_small(t, span_magic), lx = t - P * span_x;
    const int qx = ox + lx;
    float v[2][7];
#pragma unroll
    for (int h = 0; h < 2; h++) {
      const int qy = oy + 2 * P + h;
#pragma unroll
      for (int c = 0; c < 7; c++) v[h][c] = 0.f;                    // outside: depth 0 -> bilateral weight 0, colour 0
      if (qx >= 0 && qy >= 0 && qx < out.w && qy < out.h) {
        const float2 e = sample_rg16unorm_fast(normal_tex, make_tapw(texel_uv(qx, dim.w, dim.rw), texel_uv(qy, dim.h, dim.rh), normal_tex.w, normal_tex.h));
        const f3 n = decode_normal_fast(e.x, e.y);                  // uv without +0.5 (blur.comp:60,:63)
        const f3 c = unpack_rgb8unorm(ld32(refl, qx, qy));
        v[h][0] = n.x; v[h][1] = n.y; v[h][2] = n.z; v[h][3] = d24_to_float(ld32(depth1, qx, qy));
        v[h][4] = c.x; v[h][5] = c.y; v[h][6] = c.z;
      }
    }
    const int o = P * BLUR_SPAN_X + lx;
    sA[o] = make_float4(v[0][0], v[1][0], v[0][1], v[1][1]);
    sB[o] = make_float4(v[0][2], v[1][2], v[0][3], v[1][3]);
    sC[o] = make_float4(v[0][4], v[1][4], v[0][5], v[1][5]);
#if BLUR_BLUE_HALF
    sD[o] = __floats2half2_rn(v[0][6], v[1][6]);
#else
    sD[o] = make_float2(v[0][6], v[1][6]);
#endif
  }
  __syncthreads();

  float ek[2], g[2];
#pragma unroll
  for (int p = 0; p < 2; p++) {
    const int row = ty + p + R;
    const float center_depth = ((const float*)&sB[(row >> 1) * BLUR_SPAN_X + tx + R])[2 + (row & 1)];   // texelFetch(DEPTH_TEX, pixel_pos, 1)
    f3 cn = mk3(0.f, 0.f, 1.f);
    if (inside[p]) cn = sample_gbuffer_normal_fast(normal_tex, su, pixel_uv(py0 + p, dim.h, dim.rh));
    s[p].nx = dup(cn.x); s[p].ny = dup(cn.y); s[p].nz = dup(cn.z);
    s[p].ncd = dup(-center_depth);
    s[p].bk = 1000.0f * rcp_fast(center_depth);
    const float e2 = 2.f * sigma[p] * sigma[p];
    g[p] = rcp_fast(VKR_PI * e2);                                    // 1 / (2 pi sigma^2)
    ek[p] = -1.4426950408889634f * rcp_fast(e2);                     // exp(-d2/e) = exp2(d2 * ek)
    s[p].c8 = dup(exp2f(8.f * ek[p]));
    s[p].cr = s[p].cg = s[p].cb = s[p].ws = dup(0.f);
  }

  // Column walk.  The warp's window starts at staged row ty + R - wr; rounded down to a row pair (a = 0 or 1 extra row)
  // it spans np pairs.  Row h of pair k is tap j = 2k + h - wr - a - p of pixel p.
  const int row0 = ty + R - wr, a = row0 & 1, np = wr + 1 + a;
  const int base = ((row0 - a) >> 1) * BLUR_SPAN_X + tx + R;
  const ulonglong2* const pA = (const ulonglong2*)sA;
  const ulonglong2* const pB = (const ulonglong2*)sB;
  const ulonglong2* const pC = (const ulonglong2*)sC;
#if BLUR_BLUE_HALF
  const __half2* const pD = sD;
#define BLUR_LOAD(T, IDX) { (T).A = pA[IDX]; (T).B = pB[IDX]; (T).C = pC[IDX]; const float2 bb = __half22float2(pD[IDX]); (T).D = pk(bb.x, bb.y).v; }
#else
  const unsigned long long* const pD = (const unsigned long long*)sD;
#define BLUR_LOAD(T, IDX) { (T).A = pA[IDX]; (T).B = pB[IDX]; (T).C = pC[IDX]; (T).D = pD[IDX]; }
#endif
  if (__all_sync(0xFFFFFFFFu, r[0] == wr && r[1] == wr)) {
    // every pixel of the warp has the warp's radius (one material under the warp: the common case): the window is
    // the same for all lanes and only its first and last pair can hold a row outside a pixel's -r .. r
    // (row h of the first pair is inside pixel p's window iff h >= a + p, of the last pair iff h <= p - a)
    f2 gb[2], qb[2];
#pragma unroll
    for (int p = 0; p < 2; p++) {
      const int jl = -wr - a - p;
      gb[p] = pk(g[p] * exp2f((float)(jl * jl) * ek[p]), g[p] * exp2f((float)((jl + 1) * (jl + 1)) * ek[p]));
      qb[p] = pk(exp2f((float)(4 * jl + 4) * ek[p]), exp2f((float)(4 * jl + 8) * ek[p]));
    }
    const float fa = a == 0 ? 1.f : 0.f, f1 = np == 1 ? 0.f : 1.f;     // np == 1 (radius 0, aligned): the only pair is first and last
    const f2 mF[2] = {pk(fa, f1), pk(0.f, fa)}, mL[2] = {pk(fa, 0.f), pk(1.f, fa)};
    for (int i = -wr; i <= wr; i++) {
      int idx = base + i;
      const float ii = (float)(i * i);
#pragma unroll
      for (int p = 0; p < 2; p++) {
        s[p].g = mul2(gb[p], dup(exp2f(ii * ek[p])));
        s[p].q = qb[p];
      }
      BlurTap t;
      BLUR_LOAD(t, idx)
      idx += BLUR_SPAN_X;
#pragma unroll
      for (int p = 0; p < 2; p++) {
        blur_accumulate(s[p], t, mul2(mul2(s[p].g, mF[p]), blur_weight(s[p], t)));
        s[p].g = mul2(s[p].g, s[p].q); s[p].q = mul2(s[p].q, s[p].c8);
      }
#pragma unroll 2
      for (int kk = 1; kk < np - 1; kk++) {
        BLUR_LOAD(t, idx)
        idx += BLUR_SPAN_X;
#pragma unroll
        for (int p = 0; p < 2; p++) {
          blur_accumulate(s[p], t, mul2(s[p].g, blur_weight(s[p], t)));
          s[p].g = mul2(s[p].g, s[p].q); s[p].q = mul2(s[p].q, s[p].c8);
        }
      }
      if (np > 1) {
        BLUR_LOAD(t, idx)
#pragma unroll
        for (int p = 0; p < 2; p++) blur_accumulate(s[p], t, mul2(mul2(s[p].g, mL[p]), blur_weight(s[p], t)));
      }
    }
  } else {
    // mixed radii: each (pixel, row parity) runs its own recurrence from its first to its last tap inside the pixel's
    // window (pairs k0 .. k1 of the warp's walk) and carries zero outside
    int k0[2][2], k1[2][2];
    float G0[2][2], Q0[2][2], c8s[2];
#pragma unroll
    for (int p = 0; p < 2; p++) {
      unpk(s[p].c8, c8s[p], c8s[p]);
#pragma unroll
      for (int h = 0; h < 2; h++) {
        const int jb = h - wr - a - p;
        k0[p][h] = max(0, (-r[p] - jb + 1) >> 1);
        k1[p][h] = (r[p] - jb) >> 1;
        if (k0[p][h] > k1[p][h]) k0[p][h] = -1;
        const int j0 = jb + 2 * k0[p][h];
        G0[p][h] = g[p] * exp2f((float)(j0 * j0) * ek[p]);
        Q0[p][h] = exp2f((float)(4 * j0 + 4) * ek[p]);
      }
    }
    for (int i = -wr; i <= wr; i++) {
      int idx = base + i;
      const float ii = (float)(i * i);
      float gi[2], gs[2][2], qs[2][2];
#pragma unroll
      for (int p = 0; p < 2; p++) {
        gi[p] = ((i >= -r[p]) && (i <= r[p])) ? exp2f(ii * ek[p]) : 0.f;
        gs[p][0] = gs[p][1] = qs[p][0] = qs[p][1] = 0.f;
      }
      for (int kk = 0; kk < np; kk++) {
        BlurTap t;
        BLUR_LOAD(t, idx)
        idx += BLUR_SPAN_X;
#pragma unroll
        for (int p = 0; p < 2; p++) {
#pragma unroll
          for (int h = 0; h < 2; h++)
            if (kk == k0[p][h]) { gs[p][h] = G0[p][h] * gi[p]; qs[p][h] = Q0[p][h]; }
          blur_accumulate(s[p], t, mul2(pk(gs[p][0], gs[p][1]), blur_weight(s[p], t)));
#pragma unroll
          for (int h = 0; h < 2; h++) {
            gs[p][h] *= qs[p][h]; qs[p][h] *= c8s[p];
            if (kk == k1[p][h]) { gs[p][h] = 0.f; qs[p][h] = 0.f; }
          }
        }
      }
    }
  }
#undef BLUR_LOAD

#pragma unroll
  for (int p = 0; p < 2; p++) {
    if (!inside[p]) continue;
    const int py = py0 + p;
    const float inv = rcp_fast(fmaxf(hsum(s[p].ws), 0.001f));
    f3 color = mk3(hsum(s[p].cr) * inv, hsum(s[p].cg) * inv, hsum(s[p].cb) * inv);

    // temporal blend (blur.comp:77-105)
    const float sv = pixel_uv(py, dim.h, dim.rh);
    bool reprojected = false;
    const float2 vel = SAME ? unpack_rg16f(ld32(velocity, px, py)) : sample_rg16f(velocity, su, sv);
    const float delta_len = sqrt_fast(fmaf(vel.x, vel.x, vel.y * vel.y));
    const float pu = su + vel.x, pv = sv + vel.y;
    if (k.accumulate != 0 && pu >= 0 && pv >= 0 && pu <= 1 && pv <= 1) {
      const CamF cam = make_camf(k.cam);
      const f3 v_world_cur = world_pos_lod(depth1, k.inverse_camera, su, sv, cam);
      const f3 v_world_prev = world_pos_lod(prev_depth1, k.prev_inverse_camera, pu, pv, cam);
      const f3 v_camera = mk3(k.inverse_camera.m[12], k.inverse_camera.m[13], k.inverse_camera.m[14]);
      const float error = len3f(sub3f(v_world_cur, v_world_prev));
      const float pixel_dist = len3f(sub3f(v_world_cur, v_camera));
      reprojected = (delta_len < 0.0001f) || (error < fminf(fmaxf(0.1f * pixel_dist * delta_len, 0.01f), 0.1f));
    }
    if (reprojected) {
      const TapW t = make_tapw(su, sv, history.w, history.h);     // history at screen_uv, not prev_uv (App. A.6 #3)
      f3 a4 = unpack_rgb8unorm(ld32(history, t.x0, t.y0)), b4 = unpack_rgb8unorm(ld32(history, t.x1, t.y0));
      f3 c4 = unpack_rgb8unorm(ld32(history, t.x0, t.y1)), d4 = unpack_rgb8unorm(ld32(history, t.x1, t.y1));
      f3 hc = mk3(wsum(a4.x, b4.x, c4.x, d4.x, t), wsum(a4.y, b4.y, c4.y, d4.y, t), wsum(a4.z, b4.z, c4.z, d4.z, t));
      color = mk3(fmaf(0.1f, color.x - hc.x, hc.x), fmaf(0.1f, color.y - hc.y, hc.y), fmaf(0.1f, color.z - hc.z, hc.z));   // mix(history, colour, 0.1)
    }
    const uint32_t texel = pack_rgba8unorm(color.x, color.y, color.z, 0.f);
    uint32_t* const dst = (uint32_t*)(out.p + (size_t)py * out.pitch) + px;
    *dst = texel;
#pragma unroll
    for (int i = 0; i < 7; i++)                                                            // fused gather: the same texel into the peers' arrays (NVLink stores)
      if (i < po.n) *(uint32_t*)((unsigned char*)dst + po.d[i]) = texel;
  }
}

}  // namespace vkr
using namespace vkr;

static BlurK make_blur_k(const VkrBlurParams* p) {
  BlurK k;
  k.inverse_camera = make_mat(p->inverse_camera); k.prev_inverse_camera = make_mat(p->prev_inverse_camera);
  k.cam = make_cam(p->fovy_aspect_znear_zfar[0], p->fovy_aspect_znear_zfar[1], p->fovy_aspect_znear_zfar[2], p->fovy_aspect_znear_zfar[3]);
  k.max_roughness = p->max_roughness; k.accumulate = p->accumulate; k.disable_blur = p->disable_blur;
  return k;
}

extern "C" int vkr_sssr_blur_ex(const VkrBlurParams* p, const VkrImage* depth, const VkrImage* normal, const VkrImage* reflections,
                                const VkrImage* material, const VkrImage* history, const VkrImage* dvelocity, const VkrImage* prev_depth,
                                const VkrImage* out_blurred, const VkrLaunchExt* ext, void* stream) {
  if (!p || !image_ok(depth, VKR_FORMAT_D24_UNORM_S8_UINT) || depth->mip_count < 2 || !image_ok(prev_depth, VKR_FORMAT_D24_UNORM_S8_UINT) ||
      prev_depth->mip_count < 2 || !image_ok(normal, VKR_FORMAT_R16G16_UNORM) || !image_ok(reflections, VKR_FORMAT_R8G8B8A8_UNORM) ||
      !image_ok(material, VKR_FORMAT_R8G8B8A8_SRGB) || !image_ok(history, VKR_FORMAT_R8G8B8A8_UNORM) ||
      !image_ok(dvelocity, VKR_FORMAT_R16G16_SFLOAT) || !image_ok(out_blurred, VKR_FORMAT_R8G8B8A8_UNORM)) {
    set_error("sssr_blur: null argument or unexpected format");
    return VKR_ERR_INVALID_ARGUMENT;
  }
  if (reflections->width != out_blurred->width || reflections->height != out_blurred->height ||
      history->width != out_blurred->width || history->height != out_blurred->height) {
    set_error("sssr_blur: reflections / history / blurred sizes differ");
    return VKR_ERR_SIZE_MISMATCH;
  }
  ExtInfo x;
  int rc = resolve_ext(ext, out_blurred->height, BLUR_TH, &x, "sssr_blur", true);
  if (rc) return rc;
  for (int i = 0; i < x.layers; i++)
    if (!(p[i].max_roughness >= 0.f && p[i].max_roughness <= 1.f)) {   // radius bound r <= 11 assumes sigma <= 4
      set_error("sssr_blur: max_roughness outside [0,1]");
      return VKR_ERR_UNSUPPORTED;
    }
  if (history->base == out_blurred->base) { set_error("sssr_blur: history and output alias"); return VKR_ERR_INVALID_ARGUMENT; }
  const float* srgb = srgb_decode_table_device();
  if (!srgb) { set_error("sssr_blur: sRGB table upload failed"); return VKR_ERR_CUDA; }
  cudaStream_t st = (cudaStream_t)stream;
  BlurK k = make_blur_k(p);
  const BlurK* kv = nullptr;
  if (x.layers > 1) {
    if (!x.prepared()) {
      std::vector<BlurK> ks(x.layers);
      for (int i = 0; i < x.layers; i++) ks[i] = make_blur_k(p + i);
      if ((rc = upload_params(ks.data(), sizeof(BlurK), x.layers, x.scratch(), st, "sssr_blur"))) return rc;
    }
    kv = (const BlurK*)x.scratch() + x.first();
    if (x.prepare_only()) return VKR_OK;
  }
  dim3 grid((out_blurred->width + BLUR_TW - 1) / BLUR_TW, x.yblocks, x.layers);
  const bool same = same_extent(dvelocity, out_blurred) && out_blurred->width <= 8192 && out_blurred->height <= 8192;
  auto kern = same ? (kv ? sssr_blur_kernel<true, true> : sssr_blur_kernel<true, false>) : (kv ? sssr_blur_kernel<false, true> : sssr_blur_kernel<false, false>);
  {
    // opt in to > 48 KB of dynamic shared memory: a per-device, per-instantiation attribute (set once each)
    static unsigned char done[64][4];
    int dev = 0;
    cudaGetDevice(&dev);
    const int id = (same ? 2 : 0) + (kv ? 1 : 0);
    if (dev < 0 || dev >= 64 || !done[dev][id]) {
      if (cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, BLUR_SMEM_BYTES) != cudaSuccess) {
        set_error("sssr_blur: cannot reserve shared memory");
        return VKR_ERR_CUDA;
      }
      if (dev >= 0 && dev < 64) done[dev][id] = 1;
    }
  }
  kern<<<grid, 256, BLUR_SMEM_BYTES, st>>>(k, kv, LaunchX{x.yb0}, make_dim(out_blurred->width, out_blurred->height), make_mip(depth, 1, x, 0), make_mip(normal, 0, x, 1),
                             make_mip(reflections, 0, x, 2), make_mip(material, 0, x, 3), make_mip(history, 0, x, 4), make_mip(dvelocity, 0, x, 5),
                             make_mip(prev_depth, 1, x, 6), make_mip(out_blurred, 0, x, 7), srgb, x.peers());
  return check_launch("sssr_blur_kernel");
}

extern "C" int vkr_sssr_blur(const VkrBlurParams* p, const VkrImage* depth, const VkrImage* normal, const VkrImage* reflections,
                             const VkrImage* material, const VkrImage* history, const VkrImage* dvelocity, const VkrImage* prev_depth,
                             const VkrImage* out_blurred, void* stream) {
  return vkr_sssr_blur_ex(p, depth, normal, reflections, material, history, dvelocity, prev_depth, out_blurred, nullptr, stream);
}
