// taa.cu — TAA resolve: velocity reprojection, 4-neighbour history clamp, 0.1 blend,
// world-position rejection.  Replaces task "TAA" (reference src/taa.cpp:14-63,
// shaders/taa/resolve.comp:20-77).  target is RGBA16F (the image's real format, App. A.6 #1).
#include <vector>
#include "common.cuh"

#ifndef TAA_MIN_CTAS
#define TAA_MIN_CTAS 6   // resident CTAs per SM the register allocation must allow (tools/variant.sh sweeps it)
#endif
namespace vkr {

struct TaaK {
  Mat4 inverse_camera, prev_inverse_camera;
  CamK cam;
};

__device__ __forceinline__ f3 unpack_rgb16f(uint2 k) {
  float2 a = unpack_rg16f(k.x), b = unpack_rg16f(k.y);
  return mk3(a.x, a.y, b.x);
}

__device__ __forceinline__ f3 bilerp3(f3 a, f3 b, f3 c, f3 d, float wx, float wy) {
  return mk3(bilerp(a.x, b.x, c.x, d.x, wx, wy), bilerp(a.y, b.y, c.y, d.y, wx, wy), bilerp(a.z, b.z, c.z, d.z, wx, wy));
}

__device__ __forceinline__ int clampi(int v, int hi) { return min(max(v, 0), hi); }

// resolve.comp:72-77.  The depth footprint keeps the exact unfused bilinear (a silhouette texel with weight ~1e-4
// moves the reconstructed point by metres at far depths); the rest of the reconstruction is value-only.
__device__ __forceinline__ f3 world_pos(const Mip& depth, const Mat4& inv, float u, float v, const CamF& cam) {
  float d = sample_d24(depth, u, v);
  return mul_m4f(inv, reconstruct_fast(u, v, d, cam));
}

__device__ __forceinline__ f3 srgb_texel(const Mip& m, int x, int y, const float* __restrict__ srgb) {
  uint32_t k = ld32(m, x, y);
  return mk3(__ldg(srgb + (k & 255u)), __ldg(srgb + ((k >> 8) & 255u)), __ldg(srgb + ((k >> 16) & 255u)));
}

__device__ __forceinline__ f3 wsum3(f3 a, f3 b, f3 c, f3 d, const TapW& w) {
  return mk3(wsum(a.x, b.x, c.x, d.x, w), wsum(a.y, b.y, c.y, d.y, w), wsum(a.z, b.z, c.z, d.z, w));
}

// SAME: colour and velocity have the output's extent, so their bilinear footprint at the pixel's own centre is the
// texel itself (common.cuh tap_is_texel); otherwise the general footprint is evaluated.
template <bool SAME, bool LAYERED>
__global__ void __launch_bounds__(256, TAA_MIN_CTAS) taa_resolve_kernel(TaaK k0, const TaaK* __restrict__ kv, LaunchX lx, Dim dim, Mip history, Mip prev_depth, Mip depth,
                                                          Mip velocity, Mip color, Mip out, const float* __restrict__ srgb) {
  TaaK kl;
  if (LAYERED) {
    const int z = blockIdx.z;
    kl = kv[z];
    history = layer_of(history, z); prev_depth = layer_of(prev_depth, z); depth = layer_of(depth, z);
    velocity = layer_of(velocity, z); color = layer_of(color, z); out = layer_of(out, z);
  }
  const TaaK& k = LAYERED ? kl : k0;
  const int px = blockIdx.x * 32 + (threadIdx.x & 31), py = VKR_BLOCK_Y * 8 + (threadIdx.x >> 5);
  if (px >= out.w || py >= out.h) return;
  const float su = pixel_uv(px, dim.w, dim.rw), sv = pixel_uv(py, dim.h, dim.rh);
  const CamF cam = make_camf(k.cam);

  f3 current_color;
  float2 vel;
  if (SAME) {
    current_color = srgb_texel(color, px, py, srgb);
    vel = unpack_rg16f(ld32(velocity, px, py));
  } else {
    float4 cc = sample_srgb8(color, su, sv, srgb);
    current_color = mk3(cc.x, cc.y, cc.z);
    vel = sample_rg16f(velocity, su, sv);
  }
  const float delta_len = sqrt_fast(fmaf(vel.x, vel.x, vel.y * vel.y));
  bool reprojected = false;
  f3 out_color = current_color;
  const float pu = su + vel.x, pv = sv + vel.y;

  if (pu >= 0 && pv >= 0 && pu <= 1 && pv <= 1) {
    // The five bilinear footprints (centre + 4 textureOffset neighbours, :41-45) share a plus-shaped set of 12
    // texels and one pair of weights; fetch each texel once.
    const int W = history.w, H = history.h;
    float x = __fsub_rn(__fmul_rn(pu, (float)W), 0.5f), y = __fsub_rn(__fmul_rn(pv, (float)H), 0.5f);
    float fx = floorf(x), fy = floorf(y);
    Tap tw; tw.wx = x - fx; tw.wy = y - fy; tw.x0 = tw.x1 = tw.y0 = tw.y1 = 0;
    const TapW w = make_tapw(tw);
    int ix = (int)fx, iy = (int)fy;
    int xs[4] = {clampi(ix - 1, W - 1), clampi(ix, W - 1), clampi(ix + 1, W - 1), clampi(ix + 2, W - 1)};
    const unsigned char* rows[4];
#pragma unroll
    for (int j = 0; j < 4; j++) rows[j] = history.p + (size_t)clampi(iy - 1 + j, H - 1) * history.pitch;
    // texels as {r, g} pairs + b: the five footprints' red and green run on packed fp32 (same roundings per lane as wsum)
    f2 trg[4][4];
    float tb[4][4];
#pragma unroll
    for (int j = 0; j < 4; j++)
#pragma unroll
      for (int i = 0; i < 4; i++)
        if (!((i == 0 || i == 3) && (j == 0 || j == 3))) {
          const uint2 q = __ldg((const uint2*)rows[j] + xs[i]);
          const float2 rg = unpack_rg16f(q.x);
          trg[j][i] = pk(rg.x, rg.y);
          tb[j][i] = unpack_rg16f(q.y).x;
        }
    const f2 w00 = dup(w.w00), w10 = dup(w.w10), w01 = dup(w.w01), w11 = dup(w.w11);
#define TAA_FOOT(OUT, J, I)                                                                                                           \
  f3 OUT;                                                                                                                             \
  unpk(fma2(trg[J + 1][I + 1], w11, fma2(trg[J + 1][I], w01, fma2(trg[J][I + 1], w10, mul2(trg[J][I], w00)))), OUT.x, OUT.y);         \
  OUT.z = wsum(tb[J][I], tb[J][I + 1], tb[J + 1][I], tb[J + 1][I + 1], w);
    TAA_FOOT(hist, 1, 1)
    TAA_FOOT(c0, 1, 2)   // offset (+1, 0)
    TAA_FOOT(c1, 2, 1)   // offset (0, +1)
    TAA_FOOT(c2, 1, 0)   // offset (-1, 0)
    TAA_FOOT(c3, 0, 1)   // offset (0, -1)
#undef TAA_FOOT
    // min/max/clamp as single FMNMX (differs from the GLSL select form only when a history texel is NaN)
    f3 cmin = mk3(fminf(fminf(c0.x, c1.x), fminf(c2.x, c3.x)), fminf(fminf(c0.y, c1.y), fminf(c2.y, c3.y)), fminf(fminf(c0.z, c1.z), fminf(c2.z, c3.z)));
    f3 cmax = mk3(fmaxf(fmaxf(c0.x, c1.x), fmaxf(c2.x, c3.x)), fmaxf(fmaxf(c0.y, c1.y), fmaxf(c2.y, c3.y)), fmaxf(fmaxf(c0.z, c1.z), fmaxf(c2.z, c3.z)));
    hist = mk3(fminf(fmaxf(hist.x, cmin.x), cmax.x), fminf(fmaxf(hist.y, cmin.y), cmax.y), fminf(fmaxf(hist.z, cmin.z), cmax.z));

    f3 v_world_cur = world_pos(depth, k.inverse_camera, su, sv, cam);
    f3 v_world_prev = world_pos(prev_depth, k.prev_inverse_camera, pu, pv, cam);
    f3 v_camera = mk3(k.inverse_camera.m[12], k.inverse_camera.m[13], k.inverse_camera.m[14]);
    float error = len3f(sub3f(v_world_cur, v_world_prev));
    float pixel_dist = len3f(sub3f(v_world_cur, v_camera));
    reprojected = (delta_len < 0.005f) || (error < gclamp(0.1f * pixel_dist * delta_len, 0.01f, 0.2f));
    if (reprojected) out_color = mk3(fmaf(0.1f, current_color.x - hist.x, hist.x), fmaf(0.1f, current_color.y - hist.y, hist.y),
                                     fmaf(0.1f, current_color.z - hist.z, hist.z));   // mix(history, current, 0.1)
  }
  *((uint2*)(out.p + (size_t)py * out.pitch) + px) = make_uint2(pack_half2(out_color.x, out_color.y), pack_half2(out_color.z, 0.f));
}

}  // namespace vkr
using namespace vkr;

static TaaK make_taa_k(const VkrTaaParams* p) {
  TaaK k;
  k.inverse_camera = make_mat(p->inverse_camera); k.prev_inverse_camera = make_mat(p->prev_inverse_camera);
  k.cam = make_cam(p->fovy_aspect_znear_zfar[0], p->fovy_aspect_znear_zfar[1], p->fovy_aspect_znear_zfar[2], p->fovy_aspect_znear_zfar[3]);
  return k;
}

extern "C" int vkr_taa_resolve_ex(const VkrTaaParams* p, const VkrImage* history, const VkrImage* prev_depth, const VkrImage* depth,
                                  const VkrImage* velocity, const VkrImage* color, const VkrImage* out_target, const VkrLaunchExt* ext, void* stream) {
  if (!p || !image_ok(history, VKR_FORMAT_R16G16B16A16_SFLOAT) || !image_ok(prev_depth, VKR_FORMAT_D24_UNORM_S8_UINT) ||
      !image_ok(depth, VKR_FORMAT_D24_UNORM_S8_UINT) || !image_ok(velocity, VKR_FORMAT_R16G16_SFLOAT) ||
      !image_ok(color, VKR_FORMAT_R8G8B8A8_SRGB) || !image_ok(out_target, VKR_FORMAT_R16G16B16A16_SFLOAT)) {
    set_error("taa_resolve: null argument or unexpected format");
    return VKR_ERR_INVALID_ARGUMENT;
  }
  if (history->base == out_target->base) { set_error("taa_resolve: history and target alias"); return VKR_ERR_INVALID_ARGUMENT; }
  ExtInfo x;
  int rc = resolve_ext(ext, out_target->height, 8, &x, "taa_resolve");
  if (rc) return rc;
  const float* srgb = srgb_decode_table_device();
  if (!srgb) { set_error("taa_resolve: sRGB table upload failed"); return VKR_ERR_CUDA; }
  cudaStream_t st = (cudaStream_t)stream;
  TaaK k = make_taa_k(p);
  const TaaK* kv = nullptr;
  if (x.layers > 1) {
    if (!x.prepared()) {
      std::vector<TaaK> ks(x.layers);
      for (int i = 0; i < x.layers; i++) ks[i] = make_taa_k(p + i);
      if ((rc = upload_params(ks.data(), sizeof(TaaK), x.layers, x.scratch(), st, "taa_resolve"))) return rc;
    }
    kv = (const TaaK*)x.scratch() + x.first();
    if (x.prepare_only()) return VKR_OK;
  }
  dim3 grid((out_target->width + 31) / 32, x.yblocks, x.layers);
  const Dim dim = make_dim(out_target->width, out_target->height);
  const bool same = same_extent(color, out_target) && same_extent(velocity, out_target) && out_target->width <= 8192 && out_target->height <= 8192;
  auto kern = same ? (kv ? taa_resolve_kernel<true, true> : taa_resolve_kernel<true, false>) : (kv ? taa_resolve_kernel<false, true> : taa_resolve_kernel<false, false>);
  kern<<<grid, 256, 0, st>>>(k, kv, LaunchX{x.yb0}, dim, make_mip(history, 0, x, 0), make_mip(prev_depth, 0, x, 1), make_mip(depth, 0, x, 2),
                             make_mip(velocity, 0, x, 3), make_mip(color, 0, x, 4), make_mip(out_target, 0, x, 5), srgb);
  return check_launch("taa_resolve_kernel");
}

extern "C" int vkr_taa_resolve(const VkrTaaParams* p, const VkrImage* history, const VkrImage* prev_depth, const VkrImage* depth,
                               const VkrImage* velocity, const VkrImage* color, const VkrImage* out_target, void* stream) {
  return vkr_taa_resolve_ex(p, history, prev_depth, depth, velocity, color, out_target, nullptr, stream);
}
