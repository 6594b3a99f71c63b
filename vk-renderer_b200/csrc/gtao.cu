// gtao.cu — GTAO main (MIS with the SSR ray), 4x4 depth-aware filter, temporal accumulate.
//
// Replaces tasks "GTAO_main", "GTAO_filter", "GTAO_accumulate"
// (reference src/gtao.cpp:84-148,:198-239,:286-347; shaders/gtao/{main,filter,accum}.comp).
#include <vector>
#include "brdf.cuh"
#include "f32x2.cuh"

#ifndef GTAO_MIN_CTAS
#define GTAO_MIN_CTAS 5   // resident CTAs per SM the register allocation must allow (tools/variant.sh sweeps it)
#endif
namespace vkr {

struct GtaoK {
  Mat4 normal_mat;
  CamK cam;
  float angle_offset, weight_ratio;
  uint32_t use_mis, two_directions, reflections_only;
  int tw, th;  // tex_size = truncated dispatch extent (gtao.cpp:145, main.comp:54)
};

#define MAX_THIKNESS 0.1f  // main.comp:82

// main.comp:84-108 — 16 bilinear depth taps along the slice direction.
// Precision: the thickness break compares view-space z of consecutive samples, and z is ill-conditioned in d, so the
// depth footprint stays the exact unfused bilinear and z the correctly rounded quotient; the horizon cosine itself is a
// running max of a continuous quantity (value-only: rsqrt normalisation, FMA dot products).
// TILE: the CTA has staged the decoded depth texels its taps can reach (clamp-to-edge applied while staging) in shared
// memory, so a tap is 4 LDS instead of 4 clamped address computations + 4 LDG + 4 D24 decodes.  The bilinear footprint
// runs on packed fp32 (x with y, top row with bottom row); every operation rounds like the scalar unfused form.
struct DepthTile { const float* s; int ox, oy; };
#define GT_SW 72
#define GT_SH 48
template <bool TILE>
__device__ __forceinline__ float find_horizon(const Mip& depth, const DepthTile& tile, float su, float sv, f3 camera_start, float dirx, float diry,
                                              f3 v, const CamF& cam) {
  float h_cos = -1.0f;
  float previous_z = camera_start.z;
  const f2 one = one2(), suv = pk(su, sv), dir = pk(dirx, diry), wh = pk((float)depth.w, (float)depth.h);
  const f2 tg2 = pk(cam.aspect_tg, cam.tg), ncs = pk(-camera_start.x, -camera_start.y);
#pragma unroll 1
  for (int i = 1; i <= 16; i++) {
    const float s = (float)i * 0.0625f;                                      // float(i) / 16, exact
    const f2 tuv = mul_add2_unfused(dup(s), dir, suv, one);                  // FA(su, FM(s, dirx)), FA(sv, FM(s, diry))
    float d;
    if (TILE) {
      // make_tap + bilerp of common.cuh on the staged window
      const f2 xy = mul_add2_unfused(tuv, wh, dup(-0.5f), one);
      float x, y, wx, wy, omx, omy, top, bot;
      unpk(xy, x, y);
      const float fx = floorf(x), fy = floorf(y);
      const f2 w2 = sub2(xy, pk(fx, fy));
      unpk(w2, wx, wy);
      unpk(sub2(dup(1.0f), w2), omx, omy);
      int idx = ((int)fy - tile.oy) * GT_SW + ((int)fx - tile.ox);
      idx = min(max(idx, 0), GT_SW * (GT_SH - 1) - 2);                       // memory safety only: in range by construction of the window
      const float* q = tile.s + idx;
      unpk(gmix2(pk(q[0], q[GT_SW]), pk(q[1], q[GT_SW + 1]), omx, wx, one), top, bot);
      d = FA(FM(top, omy), FM(bot, wy));
    } else {
      float tu, tv;
      unpk(tuv, tu, tv);
      d = sample_d24(depth, tu, tv);
    }
    const float z = linearize_rn(d, cam);
    if (z > FA(previous_z, MAX_THIKNESS)) break;
    previous_z = z;
    float ox, oy;
    unpk(fma2(fma2(dup(-2.f), tuv, dup(1.0f)), mul2(dup(z), tg2), ncs), ox, oy);
    const f3 off = mk3(ox, oy, z - camera_start.z);
    const float sample_cos = dot3f(v, off) * rsqrt_fast(dot3f(off, off));   // no denormal fix-up: |off|^2 < 1e-38 is a zero offset
    h_cos = fmaxf(h_cos, sample_cos);                                        // a NaN (zero offset) is ignored by both max forms
  }
  return h_cos;
}

// main.comp:276-278
__device__ __forceinline__ float gtao_direction(int x, int y) { return (1.0f / 16.0f) * (float)((((x + y) & 3) << 2) + (x & 3)); }

// slice integral shared by mis_gtao (:232-246) and gtao_camera_space (:199-214).
// The projected-normal angle n = pi/2 - acos(dot(normalize(np), X)) sits on the edge of acos' domain for grazing
// surfaces: whether the argument rounds to 1 or 1 + 2^-23 decides between a finite angle and NaN, and a NaN texel
// spreads through the 4x4 filter and the temporal history.  Everything feeding it therefore keeps the oracle's
// unfused IEEE evaluation (once per pixel); the 16-tap horizon search and the closing integral are value-only.
template <bool TILE>
__device__ __forceinline__ float slice_occlusion(const Mip& depth, const DepthTile& tile, const GtaoK& k, const CamF& cam, float su, float sv, float frag_depth, f3 camera_pos,
                                                 f3 w0, f3 camera_normal, float rx, float ry, float angle, f3* Lout, bool mis) {
  const float sdx = FM(rx, cosf(angle)), sdy = FM(ry, sinf(angle));
  const f3 sample_end_pos = reconstruct_view_vec(FA(su, sdx), FA(sv, sdy), frag_depth, k.cam);
  if (Lout) *Lout = norm3(sample_end_pos - camera_pos);
  const f3 slice_normal = norm3(cross3(w0, -sample_end_pos));
  const f3 normal_projected = camera_normal - dot3(camera_normal, slice_normal) * slice_normal;
  const f3 X = -norm3(cross3(slice_normal, w0));
  const float n = FS(VKR_PI / 2.0f, acosf(dot3(norm3(normal_projected), X)));
  const float h_cos = find_horizon<TILE>(depth, tile, su, sv, camera_pos, sdx, sdy, w0, cam);
  float h = acosf(h_cos);
  h = gmin(FA(n, gmin(FS(h, n), VKR_PI / 2.0f)), h);                        // GLSL min: a NaN n propagates
  const float len = len3(normal_projected);
  const float lenp = mis ? (1.0f / VKR_PI) * len : len;                     // main.comp:246 vs :213
  return (lenp * 0.25f) * gmax(-__cosf(2.f * h - n) + __cosf(n) + 2.f * h * __sinf(n), 0.0f);
}

template <bool LAYERED>
__global__ void __launch_bounds__(256, GTAO_MIN_CTAS) gtao_main_kernel(GtaoK k0, const GtaoK* __restrict__ kv, LaunchX lx, Dim dim, Mip depth, Mip normal, Mip material, Mip pdf,
                                                        Mip raw, const float* __restrict__ srgb) {
  GtaoK kl;
  if (LAYERED) {
    const int z = blockIdx.z;
    kl = kv[z];
    depth = layer_of(depth, z); normal = layer_of(normal, z); material = layer_of(material, z); pdf = layer_of(pdf, z); raw = layer_of(raw, z);
  }
  const GtaoK& k = LAYERED ? kl : k0;
  __shared__ float s_tile[GT_SW * GT_SH];
  __shared__ int s_box[4];
  const int px = blockIdx.x * 32 + (threadIdx.x & 31), py = VKR_BLOCK_Y * 8 + (threadIdx.x >> 5);
  const bool active = px < k.tw && py < k.th;
  const float su = pixel_uv(px, dim.w, dim.rw), sv = pixel_uv(py, dim.h, dim.rh);
  const CamF cam = make_camf(k.cam);
  const float frag_depth = active ? sample_d24(depth, su, sv) : 1.f;

  // The texels this CTA's horizon taps can reach: each pixel's own texel +- its radius (<= 16 texels) plus the
  // bilinear footprint.  When the box fits the staging buffer (it does unless the depth level is much finer than the
  // AO image) the taps read decoded texels from shared memory.
  {
    int lox = 0x7fffffff, loy = 0x7fffffff, hix = -0x7fffffff, hiy = -0x7fffffff;
    if (active && frag_depth < 1.f) {
      const float rad = gmin(100.0f / len3(reconstruct_view_vec(su, sv, frag_depth, k.cam)), 16.0f);
      const float xc = fmaf(su, (float)depth.w, -0.5f), yc = fmaf(sv, (float)depth.h, -0.5f);
      if (rad >= 0.f && fabsf(xc) < 1e6f && fabsf(yc) < 1e6f) {
        lox = (int)floorf(xc - rad) - 1; hix = (int)floorf(xc + rad) + 2;
        loy = (int)floorf(yc - rad) - 1; hiy = (int)floorf(yc + rad) + 2;
      } else {
        lox = loy = -0x40000000; hix = hiy = 0x40000000;                  // not a sane camera: no staging
      }
    }
    if (threadIdx.x == 0) { s_box[0] = s_box[1] = 0x7fffffff; s_box[2] = s_box[3] = -0x7fffffff; }
    __syncthreads();
    lox = __reduce_min_sync(0xFFFFFFFFu, lox); loy = __reduce_min_sync(0xFFFFFFFFu, loy);
    hix = __reduce_max_sync(0xFFFFFFFFu, hix); hiy = __reduce_max_sync(0xFFFFFFFFu, hiy);
    if ((threadIdx.x & 31) == 0) { atomicMin(&s_box[0], lox); atomicMin(&s_box[1], loy); atomicMax(&s_box[2], hix); atomicMax(&s_box[3], hiy); }
    __syncthreads();
  }
  DepthTile tile;
  tile.s = s_tile; tile.ox = s_box[0]; tile.oy = s_box[1];
  const int bw = s_box[2] - tile.ox + 1, bh = s_box[3] - tile.oy + 1;
  const bool use_tile = s_box[2] >= tile.ox && bw <= GT_SW && bh <= GT_SH && s_box[3] >= tile.oy;
  if (use_tile) {
    const uint32_t bw_magic = div_magic((uint32_t)bw);                      // bw * bh <= GT_SW * GT_SH < 2^16
    for (int t = threadIdx.x; t < bw * bh; t += 256) {
      const int ly = div_small(t, bw_magic), lx = t - ly * bw;
      const int qx = min(max(tile.ox + lx, 0), depth.w - 1), qy = min(max(tile.oy + ly, 0), depth.h - 1);   // clamp-to-edge per index
      s_tile[ly * GT_SW + lx] = d24_to_float(ld32(depth, qx, qy));
    }
    __syncthreads();
  }
  if (!active) return;
  uint2* out = (uint2*)(raw.p + (size_t)py * raw.pitch) + px;
  float ox = 0.f, oy = 1.0f / (2 * VKR_PI);
  if (k.use_mis > 0) {                                                     // mis_gtao, main.comp:219-274
    if (frag_depth >= 1.f) { *out = make_uint2(pack_half2(0.f, 1.0f), 0u); return; }
    f3 camera_pos = reconstruct_view_vec(su, sv, frag_depth, k.cam);
    f3 w0 = -norm3(camera_pos);
    f3 camera_normal = norm3(mul_m4(k.normal_mat, sample_gbuffer_normal(normal, su, sv), 0.f));
    float rad = gmin(100.0f / len3(camera_pos), 16.0f);
    float rx = rad / (float)depth.w, ry = rad / (float)depth.h;
    float base_angle = FA(gtao_direction(px, py), k.angle_offset);
    float angle = FM(2 * VKR_PI, base_angle);
    f3 L;
    float occlusion = use_tile ? slice_occlusion<true>(depth, tile, k, cam, su, sv, frag_depth, camera_pos, w0, camera_normal, rx, ry, angle, &L, true)
                               : slice_occlusion<false>(depth, tile, k, cam, su, sv, frag_depth, camera_pos, w0, camera_normal, rx, ry, angle, &L, true);

    float roughness;
    {
      const TapW t = make_tapw(su, sv, material.w, material.h);              // only .g of the footprint is needed
#define CHG(x, y) __ldg(srgb + ((ld32(material, x, y) >> 8) & 255u))
      roughness = wsum(CHG(t.x0, t.y0), CHG(t.x1, t.y0), CHG(t.x0, t.y1), CHG(t.x1, t.y1), t);
#undef CHG
    }
    float pdf_ggx = sampleGGXdirPDF(pdf, w0, camera_normal, L, FM(roughness, roughness));
    float pdf_uniform = 1.0f / (2 * VKR_PI);
    uint2 prev = *out;
    float2 ao = unpack_rg16f(prev.x);                                       // imageLoad(gtao_out).xy
    if (k.reflections_only != 0) {
      float res = ao.x / ao.y;
      ox = isnan(res) ? 1.f : res; oy = 1.f;
    } else {
      float alpha = 1 / FA(k.weight_ratio, 1.f);
      float betta = FS(1.f, alpha);
      float mis_weight1 = alpha / FA(FM(alpha, ao.y), FM(betta, pdf_uniform));
      float mis_weight2 = betta / FA(FM(alpha, pdf_ggx), FM(betta, pdf_uniform));
      float mis_ao = FA(FM(ao.x, mis_weight1), FM(occlusion, mis_weight2));
      ox = isnan(mis_ao) ? occlusion / pdf_uniform : mis_ao; oy = 1.f;
    }
  } else {                                                                  // gtao_camera_space, main.comp:185-217
    if (frag_depth >= 1.f) ox = 0.f;
    else {
      f3 camera_pos = reconstruct_view_vec(su, sv, frag_depth, k.cam);
      f3 w0 = -norm3(camera_pos);
      f3 camera_normal = norm3(mul_m4(k.normal_mat, sample_gbuffer_normal(normal, su, sv), 0.f));
      float rad = gmin(100.0f / len3(camera_pos), 16.0f);
      float rx = rad / (float)depth.w, ry = rad / (float)depth.h;
      float base_angle = gtao_direction(px, py) + k.angle_offset;
      int dirs = k.two_directions != 0 ? 2 : 1;
      float sum = 0.f;
      for (int d = 0; d < dirs; d++) {
        float angle = 2 * VKR_PI * (base_angle + (float)d / (float)dirs);
        sum += use_tile ? slice_occlusion<true>(depth, tile, k, cam, su, sv, frag_depth, camera_pos, w0, camera_normal, rx, ry, angle, nullptr, false)
                        : slice_occlusion<false>(depth, tile, k, cam, su, sv, frag_depth, camera_pos, w0, camera_normal, rx, ry, angle, nullptr, false);
      }
      ox = 2 * sum / (float)dirs;
    }
  }
  *out = make_uint2(pack_half2(ox, oy), 0u);
}

// filter.comp:17-52 — 4x4 window at offsets -2..+1, weight on linear depth.  A 32x8 tile stages the linearised
// depth and raw.r of its 35x11 footprint once in shared memory (the shader linearises each of the 16 taps again).
#define GF_W 32
#define GF_H 8
#define GF_SW (GF_W + 3)
#define GF_SH (GF_H + 3)
__global__ void __launch_bounds__(256) gtao_filter_kernel(LaunchX lx, Mip depth, Mip raw, Mip filtered, int tw, int th, float znear, float zfar) {
  __shared__ float s_z[GF_SH][GF_SW + 1];
  __shared__ float s_r[GF_SH][GF_SW + 1];
  depth = layer_of(depth, blockIdx.z); raw = layer_of(raw, blockIdx.z); filtered = layer_of(filtered, blockIdx.z);   // znear/zfar: one set per launch
  const int bx = blockIdx.x * GF_W - 2, by = VKR_BLOCK_Y * GF_H - 2;
  const float nf = FM(znear, zfar), fmn = FS(zfar, znear);
  for (int t = threadIdx.x; t < GF_SW * GF_SH; t += 256) {
    const int ly = t / GF_SW, lx = t - ly * GF_SW;
    const int sx = bx + lx, sy = by + ly;
    const float d = fetch_d24(depth, sx, sy);                                // out of range = 0 (App. A.3)
    s_z[ly][lx] = nf * rcp_fast(FS(FM(d, fmn), zfar));                       // linearize_depth2, exact denominator
    s_r[ly][lx] = in_range(raw, sx, sy) ? __half2float(__ushort_as_half((unsigned short)(__ldg((const uint32_t*)(raw.p + (size_t)sy * raw.pitch) + 2 * sx) & 0xFFFFu))) : 0.f;
  }
  __syncthreads();
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
  const int px = blockIdx.x * GF_W + tx, py = VKR_BLOCK_Y * GF_H + ty;
  if (px >= tw || py >= th) return;
  const float linear_depth = s_z[ty + 2][tx + 2];
  const float k5 = 5.f * rcp_fast(fabsf(linear_depth));
  float weight_sum = 0, ao = 0;
#pragma unroll
  for (int x = 0; x < 4; x++) {
#pragma unroll
    for (int y = 0; y < 4; y++) {
      const float weight = fmaxf(0.0f, fmaf(-k5, fabsf(s_z[ty + y][tx + x] - linear_depth), 1.0f));
      weight_sum += weight;
      ao = fmaf(weight, s_r[ty + y][tx + x], ao);
    }
  }
  *((__half*)(filtered.p + (size_t)py * filtered.pitch) + px) = __float2half_rn(ao * rcp_fast(weight_sum));
}

struct AccumK {
  Mat4 inverse_camera, prev_inverse_camera, mvp;
  CamK cam;
  uint32_t clear_history;
};

// accum.comp:29-99.  Depth footprints keep the exact unfused bilinear and the exact z denominator (far depths amplify a
// 1-ulp difference to ~1e-3 view units, and depth_err is thresholded); everything else is value-only.
template <bool SAME, bool LAYERED>
__global__ void __launch_bounds__(256) gtao_accumulate_kernel(AccumK k0, const AccumK* __restrict__ kv, LaunchX lx, Dim dim, Mip depth, Mip prev_depth, Mip filtered,
                                                              Mip out, Mip velocity, Mip history, PeerOut po) {
  AccumK kl;
  if (LAYERED) {
    const int z = blockIdx.z;
    kl = kv[z];
    depth = layer_of(depth, z); prev_depth = layer_of(prev_depth, z); filtered = layer_of(filtered, z); out = layer_of(out, z);
    velocity = layer_of(velocity, z); history = layer_of(history, z);
  }
  const AccumK& k = LAYERED ? kl : k0;
  const int px = blockIdx.x * 32 + (threadIdx.x & 31), py = VKR_BLOCK_Y * 8 + (threadIdx.x >> 5);
  if (px >= out.w || py >= out.h) return;
  const float tw = dim.w, th = dim.h;
  const float su = pixel_uv(px, dim.w, dim.rw), sv = pixel_uv(py, dim.h, dim.rh);
  const CamF cam = make_camf(k.cam);
  const float2 vel = SAME ? unpack_rg16f(ld32(velocity, px, py)) : sample_rg16f(velocity, su, sv);
  const float pu = FA(su, vel.x), pv = FA(sv, vel.y);
  bool reprojected = false;
  float valid_samples = 1.0f;
  if (pu >= 0 && pv >= 0 && pu <= 1 && pv <= 1) {
    // accum.comp:51-71.  `reprojected` is a discontinuous choice whose inputs (depth, prev_depth, velocity, camera) are bit-identical
    // on both sides of a parity run, and the running average it gates has a 1/N memory that the bilinear history fetch then
    // spreads: a texel that flips here is still visible -- and wider -- 50 frames later (tests/test_parity_gpu.py, 64-frame
    // chain: 3 % of the texels beyond 2e-3 at frame 63 with the value-only arithmetic this block used to have).  So the whole
    // decision path carries the oracle's bits: unfused products and sums in the shader's order, correctly rounded quotients.
    const float d_prev = sample_d24(prev_depth, pu, pv);
    const float z_cam = linearize_rn(d_prev, cam);                                                 // gbuffer_encode.glsl:53-69
    const f3 v_camera = mk3(FM(-FS(FM(2.f, pu), 1.f), FM(FM(z_cam, k.cam.aspect), k.cam.tg)), FM(-FS(FM(2.f, pv), 1.f), FM(z_cam, k.cam.tg)), z_cam);
    const f3 v_world_prev = mul_m4(k.prev_inverse_camera, v_camera, 1.0f);
    const float4 clip = mul_m4_4(k.mvp, v_world_prev, 1.0f);
    const float ndx = div_rn_normal(clip.x, clip.w), ndy = div_rn_normal(clip.y, clip.w), ndz = div_rn_normal(clip.z, clip.w);
    const float wu = FA(FM(0.5f, ndx), 0.5f), wv = FA(FM(0.5f, ndy), 0.5f);
    const float dx = FM(fabsf(FS(wu, su)), tw), dy = FM(fabsf(FS(wv, sv)), th);
    const float current_z = linearize_rn(sample_d24(depth, su, sv), cam);
    const float prev_z = linearize_rn(ndz, cam);
    const float depth_err = fabsf(FS(prev_z, current_z));
    const float vel_delta = gmax(FM(fabsf(vel.x), tw), FM(fabsf(vel.y), th));
    const float error = FA(FM(0.1f, vel_delta), depth_err);
    valid_samples = fminf(fmaxf(1.f - error, 0.8f), 1.f);
    reprojected = (gmax(dx, dy) <= 2.0f) && (depth_err < 0.2f);
  }
  const float new_ao = __half2float(*((const __half*)(filtered.p + (size_t)py * filtered.pitch) + px));
  float computed_ao = new_ao;
  float samples_count = 1.f;
  if (k.clear_history != 0) reprojected = false;
  if (reprojected) {
    const float2 acc = sample_rg16f_fast(history, make_tapw(pu, pv, history.w, history.h));
    samples_count = 255.f * acc.y * valid_samples;
    computed_ao = fmaf(acc.x, samples_count, new_ao) * rcp_fast(samples_count + 1.f);
    samples_count = samples_count + 1.f;
    if (samples_count > 255.f) samples_count = 100.f;
  }
  const uint32_t texel = pack_half2(gclamp(computed_ao, 0.f, 1.f), samples_count * (1.0f / 255.f));
  uint32_t* const dst = (uint32_t*)(out.p + (size_t)py * out.pitch) + px;
  *dst = texel;
#pragma unroll
  for (int i = 0; i < 7; i++)                                                              // fused gather: the same texel into the peers' arrays (NVLink stores)
    if (i < po.n) *(uint32_t*)((unsigned char*)dst + po.d[i]) = texel;
}

}  // namespace vkr
using namespace vkr;

static GtaoK make_gtao_k(const VkrGtaoParams* p, const VkrImage* raw) {
  GtaoK k;
  k.normal_mat = make_mat(p->normal_mat);
  k.cam = make_cam(p->fovy, p->aspect, p->znear, p->zfar);
  k.angle_offset = p->angle_offset; k.weight_ratio = p->weight_ratio;
  k.use_mis = p->use_mis; k.two_directions = p->two_directions; k.reflections_only = p->reflections_only;
  k.tw = (int)(raw->width / 8) * 8; k.th = (int)(raw->height / 4) * 4;  // truncating dispatch, gtao.cpp:145
  return k;
}

extern "C" int vkr_gtao_main_ex(const VkrGtaoParams* p, const VkrImage* depth, const VkrImage* normal, const VkrImage* material,
                                const VkrImage* pdf_lut, const VkrImage* raw, uint32_t depth_lod, const VkrLaunchExt* ext, void* stream) {
  if (!p || !image_ok(depth, VKR_FORMAT_D24_UNORM_S8_UINT) || depth_lod >= depth->mip_count || !image_ok(normal, VKR_FORMAT_R16G16_UNORM) ||
      !image_ok(material, VKR_FORMAT_R8G8B8A8_SRGB) || !image_ok(pdf_lut, VKR_FORMAT_R32_SFLOAT) || !image_ok(raw, VKR_FORMAT_R16G16B16A16_SFLOAT)) {
    set_error("gtao_main: null argument or unexpected format");
    return VKR_ERR_INVALID_ARGUMENT;
  }
  GtaoK k = make_gtao_k(p, raw);
  if (k.tw == 0 || k.th == 0) return VKR_OK;
  ExtInfo x;
  int rc = resolve_ext(ext, (uint32_t)k.th, 8, &x, "gtao_main");
  if (rc) return rc;
  const float* srgb = srgb_decode_table_device();
  if (!srgb) { set_error("gtao_main: sRGB table upload failed"); return VKR_ERR_CUDA; }
  cudaStream_t st = (cudaStream_t)stream;
  const GtaoK* kv = nullptr;
  if (x.layers > 1) {
    if (!x.prepared()) {
      std::vector<GtaoK> ks(x.layers);
      for (int i = 0; i < x.layers; i++) ks[i] = make_gtao_k(p + i, raw);
      if ((rc = upload_params(ks.data(), sizeof(GtaoK), x.layers, x.scratch(), st, "gtao_main"))) return rc;
    }
    kv = (const GtaoK*)x.scratch() + x.first();
    if (x.prepare_only()) return VKR_OK;
  }
  dim3 grid((k.tw + 31) / 32, x.yblocks, x.layers);
  auto kern = kv ? gtao_main_kernel<true> : gtao_main_kernel<false>;
  kern<<<grid, 256, 0, st>>>(k, kv, LaunchX{x.yb0}, make_dim(k.tw, k.th), make_mip(depth, depth_lod, x, 0), make_mip(normal, 0, x, 1), make_mip(material, 0, x, 2),
                             make_mip(pdf_lut, 0, x, 3), make_mip(raw, 0, x, 4), srgb);
  return check_launch("gtao_main_kernel");
}

extern "C" int vkr_gtao_main(const VkrGtaoParams* p, const VkrImage* depth, const VkrImage* normal, const VkrImage* material,
                             const VkrImage* pdf_lut, const VkrImage* raw, uint32_t depth_lod, void* stream) {
  return vkr_gtao_main_ex(p, depth, normal, material, pdf_lut, raw, depth_lod, nullptr, stream);
}

extern "C" int vkr_gtao_filter_ex(const VkrGtaoParams* p, const VkrImage* depth, const VkrImage* raw, const VkrImage* filtered,
                                  uint32_t depth_lod, const VkrLaunchExt* ext, void* stream) {
  if (!p || !image_ok(depth, VKR_FORMAT_D24_UNORM_S8_UINT) || depth_lod >= depth->mip_count ||
      !image_ok(raw, VKR_FORMAT_R16G16B16A16_SFLOAT) || !image_ok(filtered, VKR_FORMAT_R16_SFLOAT)) {
    set_error("gtao_filter: null argument or unexpected format");
    return VKR_ERR_INVALID_ARGUMENT;
  }
  int tw = (int)(filtered->width / 8) * 8, th = (int)(filtered->height / 4) * 4;  // gtao.cpp:237
  if (tw == 0 || th == 0) return VKR_OK;
  ExtInfo x;
  int rc = resolve_ext(ext, (uint32_t)th, 8, &x, "gtao_filter");
  if (rc) return rc;
  for (int i = 1; i < x.layers; i++)
    if (p[i].znear != p[0].znear || p[i].zfar != p[0].zfar) { set_error("gtao_filter: layers must share znear/zfar"); return VKR_ERR_UNSUPPORTED; }
  if (x.prepare_only()) return VKR_OK;
  dim3 grid((tw + 31) / 32, x.yblocks, x.layers);
  gtao_filter_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(LaunchX{x.yb0}, make_mip(depth, depth_lod, x, 0), make_mip(raw, 0, x, 1), make_mip(filtered, 0, x, 2), tw, th,
                                                             p->znear, p->zfar);
  return check_launch("gtao_filter_kernel");
}

extern "C" int vkr_gtao_filter(const VkrGtaoParams* p, const VkrImage* depth, const VkrImage* raw, const VkrImage* filtered,
                               uint32_t depth_lod, void* stream) {
  return vkr_gtao_filter_ex(p, depth, raw, filtered, depth_lod, nullptr, stream);
}

static AccumK make_accum_k(const VkrAccumParams* p) {
  AccumK k;
  k.inverse_camera = make_mat(p->inverse_camera); k.prev_inverse_camera = make_mat(p->prev_inverse_camera); k.mvp = make_mat(p->mvp);
  k.cam = make_cam(p->fovy_aspect_znear_zfar[0], p->fovy_aspect_znear_zfar[1], p->fovy_aspect_znear_zfar[2], p->fovy_aspect_znear_zfar[3]);
  k.clear_history = p->clear_history;
  return k;
}

extern "C" int vkr_gtao_accumulate_ex(const VkrAccumParams* p, const VkrImage* depth, const VkrImage* prev_depth, const VkrImage* filtered,
                                      const VkrImage* accumulated, const VkrImage* dvelocity, const VkrImage* history, uint32_t depth_lod,
                                      const VkrLaunchExt* ext, void* stream) {
  if (!p || !image_ok(depth, VKR_FORMAT_D24_UNORM_S8_UINT) || !image_ok(prev_depth, VKR_FORMAT_D24_UNORM_S8_UINT) ||
      depth_lod >= depth->mip_count || depth_lod >= prev_depth->mip_count || !image_ok(filtered, VKR_FORMAT_R16_SFLOAT) ||
      !image_ok(accumulated, VKR_FORMAT_R16G16_SFLOAT) || !image_ok(dvelocity, VKR_FORMAT_R16G16_SFLOAT) || !image_ok(history, VKR_FORMAT_R16G16_SFLOAT)) {
    set_error("gtao_accumulate: null argument or unexpected format");
    return VKR_ERR_INVALID_ARGUMENT;
  }
  if (filtered->width != accumulated->width || filtered->height != accumulated->height) {
    set_error("gtao_accumulate: filtered and accumulated sizes differ");
    return VKR_ERR_SIZE_MISMATCH;
  }
  ExtInfo x;
  int rc = resolve_ext(ext, accumulated->height, 8, &x, "gtao_accumulate", true);
  if (rc) return rc;
  cudaStream_t st = (cudaStream_t)stream;
  AccumK k = make_accum_k(p);
  const AccumK* kv = nullptr;
  if (x.layers > 1) {
    if (!x.prepared()) {
      std::vector<AccumK> ks(x.layers);
      for (int i = 0; i < x.layers; i++) ks[i] = make_accum_k(p + i);
      if ((rc = upload_params(ks.data(), sizeof(AccumK), x.layers, x.scratch(), st, "gtao_accumulate"))) return rc;
    }
    kv = (const AccumK*)x.scratch() + x.first();
    if (x.prepare_only()) return VKR_OK;
  }
  dim3 grid((accumulated->width + 31) / 32, x.yblocks, x.layers);
  const Dim dim = make_dim(accumulated->width, accumulated->height);
  const bool same = same_extent(dvelocity, accumulated) && accumulated->width <= 8192 && accumulated->height <= 8192;
  auto kern = same ? (kv ? gtao_accumulate_kernel<true, true> : gtao_accumulate_kernel<true, false>)
                   : (kv ? gtao_accumulate_kernel<false, true> : gtao_accumulate_kernel<false, false>);
  kern<<<grid, 256, 0, st>>>(k, kv, LaunchX{x.yb0}, dim, make_mip(depth, depth_lod, x, 0), make_mip(prev_depth, depth_lod, x, 1), make_mip(filtered, 0, x, 2),
                             make_mip(accumulated, 0, x, 3), make_mip(dvelocity, 0, x, 4), make_mip(history, 0, x, 5), x.peers());
  return check_launch("gtao_accumulate_kernel");
}

extern "C" int vkr_gtao_accumulate(const VkrAccumParams* p, const VkrImage* depth, const VkrImage* prev_depth, const VkrImage* filtered,
                                   const VkrImage* accumulated, const VkrImage* dvelocity, const VkrImage* history, uint32_t depth_lod, void* stream) {
  return vkr_gtao_accumulate_ex(p, depth, prev_depth, filtered, accumulated, dvelocity, history, depth_lod, nullptr, stream);
}
