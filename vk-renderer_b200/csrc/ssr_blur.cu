// ssr_blur.cu — roughness-sized bilateral Gaussian blur of the resolved reflections plus the
// temporal blend.  Replaces task "SSSR_blur" (reference src/advanced_ssr.cpp:371-438,
// shaders/advanced_ssr/blur.comp:31-118).
//
// The reference shader fetches, for each of up to (2*11+1)^2 = 529 taps per pixel, one depth
// texel, one bilinear normal (4 texels + octahedral decode + normalize) and one colour texel.
// All three depend only on the tap POSITION, so a CTA of 16x16 pixels stages them once per
// position for its tile + halo (halo = the tile's largest radius) in shared memory as two
// float4 planes {n.xyz, depth} and {rgb, -}; the tap loop is then 2 LDS.128 + ~16 FP ops.
// The bilateral weights are per tap, so the filter is not separable.
// Warps iterate the window of their largest radius in lockstep (lanes outside their own
// radius skip), which keeps the shared-memory reads conflict-free and preserves each pixel's
// accumulation order (i = x offset outer, j = y offset inner, as in blur.comp:57-75).
#include <vector>
#include "common.cuh"

namespace vkr {

struct BlurK {
  Mat4 inverse_camera, prev_inverse_camera;
  CamK cam;
  float max_roughness;
  uint32_t accumulate, disable_blur;
};

#define BLUR_TILE 16
#define BLUR_RMAX 11
#define BLUR_SPAN (BLUR_TILE + 2 * BLUR_RMAX)

// blur.comp:110-115 — exact depth footprint and z denominator (thresholded downstream), value-only otherwise
__device__ __forceinline__ f3 world_pos_lod(const Mip& depth1, const Mat4& inv, float u, float v, const CamF& cam) {
  return mul_m4f(inv, reconstruct_fast(u, v, sample_d24(depth1, u, v), cam));
}

template <bool SAME, bool LAYERED>
__global__ void __launch_bounds__(256) sssr_blur_kernel(BlurK k0, const BlurK* __restrict__ kv, LaunchX lx, Dim dim, Mip depth1, Mip normal_tex, Mip refl,
                                                        Mip material_tex, Mip history, Mip velocity, Mip prev_depth1, Mip out, const float* __restrict__ srgb) {
  __shared__ float4 s_nd[BLUR_SPAN * BLUR_SPAN];
  __shared__ float4 s_col[BLUR_SPAN * BLUR_SPAN];
  __shared__ int s_rmax;
  BlurK kl;
  if (LAYERED) {
    const int z = blockIdx.z;
    kl = kv[z];
    depth1 = layer_of(depth1, z); normal_tex = layer_of(normal_tex, z); refl = layer_of(refl, z); material_tex = layer_of(material_tex, z);
    history = layer_of(history, z); velocity = layer_of(velocity, z); prev_depth1 = layer_of(prev_depth1, z); out = layer_of(out, z);
  }
  const BlurK& k = LAYERED ? kl : k0;

  const int tx = threadIdx.x & 15, ty = threadIdx.x >> 4;
  const int px = blockIdx.x * BLUR_TILE + tx, py = VKR_BLOCK_Y * BLUR_TILE + ty;
  const bool inside = px < out.w && py < out.h;
  const float su = pixel_uv(px, dim.w, dim.rw), sv = pixel_uv(py, dim.h, dim.rh);

  // per-pixel radius (blur.comp:39-53).  floor() makes this the one discontinuous quantity of the pass: unfused.
  float sigma = 0.35f;
  int r = 0;
  if (inside) {
    float roughness = sample_srgb8(material_tex, su, sv, srgb).y;
    roughness = gmix(0.0f, k.max_roughness, roughness);
    sigma = gmix(0.4f, 4.f, roughness);
    if (k.disable_blur != 0) sigma = 0.35f;
    r = (int)floorf(FS(FM(3.f, sigma), 0.01f));
    r = min(max(r, 0), BLUR_RMAX);
  }
  if (threadIdx.x == 0) s_rmax = 0;
  __syncthreads();
  const int wr = __reduce_max_sync(0xFFFFFFFFu, r);
  if ((threadIdx.x & 31) == 0) atomicMax(&s_rmax, wr);
  __syncthreads();
  const int R = s_rmax;
  const int span = BLUR_TILE + 2 * R;

  // stage {normal, depth} and colour for every tap position of the tile
  const int ox = blockIdx.x * BLUR_TILE - R, oy = VKR_BLOCK_Y * BLUR_TILE - R;
  for (int t = threadIdx.x; t < span * span; t += 256) {
    const int ly = t / span, lx = t - ly * span;
    const int qx = ox + lx, qy = oy + ly;
    float4 nd = make_float4(0.f, 0.f, 0.f, 0.f), col = make_float4(0.f, 0.f, 0.f, 0.f);
    if (qx >= 0 && qy >= 0 && qx < out.w && qy < out.h) {          // outside: depth 0 -> bilateral weight 0, colour 0
      const float2 e = sample_rg16unorm_fast(normal_tex, make_tapw(texel_uv(qx, dim.w, dim.rw), texel_uv(qy, dim.h, dim.rh), normal_tex.w, normal_tex.h));
      const f3 n = decode_normal_fast(e.x, e.y);                    // uv without +0.5 (blur.comp:60,:63)
      nd = make_float4(n.x, n.y, n.z, d24_to_float(ld32(depth1, qx, qy)));
      const f3 c = unpack_rgb8unorm(ld32(refl, qx, qy));
      col = make_float4(c.x, c.y, c.z, 0.f);
    }
    s_nd[ly * BLUR_SPAN + lx] = nd;
    s_col[ly * BLUR_SPAN + lx] = col;
  }
  __syncthreads();

  const float center_depth = s_nd[(ty + R) * BLUR_SPAN + tx + R].w;            // texelFetch(DEPTH_TEX, pixel_pos, 1)
  f3 center_normal = mk3(0.f, 0.f, 1.f);
  if (inside) center_normal = sample_gbuffer_normal_fast(normal_tex, su, sv);
  const float e2 = 2.f * sigma * sigma;
  const float g = rcp_fast(VKR_PI * e2);                                       // 1 / (2 pi sigma^2)
  const float ek = -1.4426950408889634f * rcp_fast(e2);                        // exp(-d2/e) = exp2(d2 * ek)
  const float bk = 1000.0f * rcp_fast(center_depth);
  // Gaussian along a column by recurrence: G(j+1) = G(j) * q(j), q(j+1) = q(j) * c, started at the lane's own -r
  // (two multiplies per tap instead of int->float, multiply, MUFU.EX2 with range fix-up).
  const float gc = exp2f(2.f * ek);
  const float q0 = exp2f((float)(1 - 2 * r) * ek);
  const float g0 = g * exp2f((float)(r * r) * ek);

  float weight_sum = 0.f;
  float cr = 0.f, cg = 0.f, cb = 0.f;
  const float4* pn0 = s_nd + (ty + R - wr) * BLUR_SPAN + tx + R;
  const float4* pc0 = s_col + (ty + R - wr) * BLUR_SPAN + tx + R;
#define BLUR_TAP(GJ)                                                                                                              \
  {                                                                                                                               \
    const float4 nd = *pn;                                                                                                        \
    const float4 c = *pc;                                                                                                         \
    pn += BLUR_SPAN; pc += BLUR_SPAN;                                                                                             \
    const float bilateral_weight = fmaxf(fmaf(-fabsf(center_depth - nd.w), bk, 1.0f), 0.f);                                       \
    const float normal_weight = fmaxf(fmaf(center_normal.z, nd.z, fmaf(center_normal.y, nd.y, center_normal.x * nd.x)), 0.f);     \
    const float w = (GJ) * (bilateral_weight * normal_weight);                                                                    \
    cr = fmaf(c.x, w, cr); cg = fmaf(c.y, w, cg); cb = fmaf(c.z, w, cb);                                                          \
    weight_sum += w;                                                                                                              \
  }
  if (__all_sync(0xFFFFFFFFu, r == wr)) {
    // every lane of the warp has the warp's radius (one material under the warp: the common case): no per-tap predicates
    for (int i = -wr; i <= wr; i++) {
      const float4* pn = pn0 + i;
      const float4* pc = pc0 + i;
      float gj = g0 * exp2f((float)(i * i) * ek), q = q0;
#pragma unroll 4
      for (int j = -wr; j <= wr; j++) {
        BLUR_TAP(gj)
        gj *= q; q *= gc;
      }
    }
  } else {
    // mixed radii: lanes outside their own window carry a zero Gaussian (gj = 0 before -r and after +r)
    for (int i = -wr; i <= wr; i++) {
      const float gi = ((i >= -r) && (i <= r)) ? exp2f((float)(i * i) * ek) : 0.f;
      const float4* pn = pn0 + i;
      const float4* pc = pc0 + i;
      float gj = 0.f, q = 0.f;
      for (int j = -wr; j <= wr; j++) {
        if (j == -r) { gj = g0 * gi; q = q0; }
        BLUR_TAP(gj)
        gj *= q; q *= gc;
        if (j == r) { gj = 0.f; q = 0.f; }
      }
    }
  }
#undef BLUR_TAP
  if (!inside) return;
  const float inv = rcp_fast(fmaxf(weight_sum, 0.001f));
  f3 color = mk3(cr * inv, cg * inv, cb * inv);

  // temporal blend (blur.comp:77-105)
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
    f3 a = unpack_rgb8unorm(ld32(history, t.x0, t.y0)), b = unpack_rgb8unorm(ld32(history, t.x1, t.y0));
    f3 c = unpack_rgb8unorm(ld32(history, t.x0, t.y1)), d = unpack_rgb8unorm(ld32(history, t.x1, t.y1));
    f3 hc = mk3(wsum(a.x, b.x, c.x, d.x, t), wsum(a.y, b.y, c.y, d.y, t), wsum(a.z, b.z, c.z, d.z, t));
    color = mk3(fmaf(0.1f, color.x - hc.x, hc.x), fmaf(0.1f, color.y - hc.y, hc.y), fmaf(0.1f, color.z - hc.z, hc.z));   // mix(history, colour, 0.1)
  }
  *((uint32_t*)(out.p + (size_t)py * out.pitch) + px) = pack_rgba8unorm(color.x, color.y, color.z, 0.f);
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
  int rc = resolve_ext(ext, out_blurred->height, BLUR_TILE, &x, "sssr_blur");
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
    kv = (const BlurK*)x.scratch();
    if (x.prepare_only()) return VKR_OK;
  }
  dim3 grid((out_blurred->width + BLUR_TILE - 1) / BLUR_TILE, x.yblocks, x.layers);
  const bool same = same_extent(dvelocity, out_blurred) && out_blurred->width <= 8192 && out_blurred->height <= 8192;
  auto kern = same ? (kv ? sssr_blur_kernel<true, true> : sssr_blur_kernel<true, false>) : (kv ? sssr_blur_kernel<false, true> : sssr_blur_kernel<false, false>);
  kern<<<grid, 256, 0, st>>>(k, kv, LaunchX{x.yb0}, make_dim(out_blurred->width, out_blurred->height), make_mip(depth, 1, x, 0), make_mip(normal, 0, x, 1),
                             make_mip(reflections, 0, x, 2), make_mip(material, 0, x, 3), make_mip(history, 0, x, 4), make_mip(dvelocity, 0, x, 5),
                             make_mip(prev_depth, 1, x, 6), make_mip(out_blurred, 0, x, 7), srgb);
  return check_launch("sssr_blur_kernel");
}

extern "C" int vkr_sssr_blur(const VkrBlurParams* p, const VkrImage* depth, const VkrImage* normal, const VkrImage* reflections,
                             const VkrImage* material, const VkrImage* history, const VkrImage* dvelocity, const VkrImage* prev_depth,
                             const VkrImage* out_blurred, void* stream) {
  return vkr_sssr_blur_ex(p, depth, normal, reflections, material, history, dvelocity, prev_depth, out_blurred, nullptr, stream);
}
