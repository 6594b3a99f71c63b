// ssr_filter.cu — 5-tap BRDF-weighted resolve of the traced rays.
// Replaces task "SSSR_filter" (reference src/advanced_ssr.cpp:308-369,
// shaders/advanced_ssr/filter.comp:36-149).  Quirks kept on purpose:
// uv = pixel / size without +0.5 (#5) and the swapped brdfG1 arguments (#4).
#include <vector>
#include "brdf.cuh"

#ifndef FILTER_MIN_CTAS
#define FILTER_MIN_CTAS 6   // resident CTAs per SM the register allocation must allow (tools/variant.sh sweeps it)
#endif
namespace vkr {

struct FilterK {
  Mat4 normal_mat;
  CamK cam;
  uint32_t render_flags;
};

// Everything process_pixel() (filter.comp:110-149) derives from the TAP position alone — the ray, its
// reconstructed hit, the radiance fetched at the hit, V, L, H, N.L, N.V and the Fresnel factor — is
// computed once per position of a (32+2) x (8+2) tile and staged in shared memory; a pixel's five taps
// then only combine those with its own F0 / roughness / centre depth.  (The reference evaluates the
// whole of process_pixel five times per pixel.)  Out-of-range positions are processed like the
// reference does: texelFetch returns zeros and the arithmetic runs on them.
struct TapRec {
  float fs, tl, tv, ndv;      // (1 - H.V)^5 ; (1-NdotL^2)/NdotL^2 ; (1-NdotV^2)/NdotV^2 ; NdotV
  float r, g, b, depth;       // radiance at the hit ; mip-1 depth of the tap pixel
};

#define FT_W 32
#define FT_H 8

// Nothing here selects between discrete alternatives (ray validity is an exact unorm16 compare), so the whole pass runs on
// the value-only helpers of common.cuh; depth linearisation keeps its exact denominator.
template <bool LAYERED>
__global__ void __launch_bounds__(256, FILTER_MIN_CTAS) sssr_filter_kernel(FilterK k0, const FilterK* __restrict__ kv, LaunchX lx, Dim dim, Mip rays, Mip depth1, Mip albedo_tex,
                                                          Mip normal_tex, Mip material_tex, Mip out, const float* __restrict__ srgb) {
  __shared__ float4 s_a[(FT_W + 2) * (FT_H + 2)];
  __shared__ float4 s_b[(FT_W + 2) * (FT_H + 2)];
  FilterK kl;
  if (LAYERED) {
    const int z = blockIdx.z;
    kl = kv[z];
    rays = layer_of(rays, z); depth1 = layer_of(depth1, z); albedo_tex = layer_of(albedo_tex, z); normal_tex = layer_of(normal_tex, z);
    material_tex = layer_of(material_tex, z); out = layer_of(out, z);
  }
  const FilterK& k = LAYERED ? kl : k0;
  const int bx = blockIdx.x * FT_W - 1, by = VKR_BLOCK_Y * FT_H - 1;
  const CamF cam = make_camf(k.cam);
#define CH(k, s) __ldg(srgb + (((k) >> (s)) & 255u))

  for (int t = threadIdx.x; t < (FT_W + 2) * (FT_H + 2); t += 256) {
    const int ly = t / (FT_W + 2), lx = t - ly * (FT_W + 2);
    const int qx = bx + lx, qy = by + ly;
    float4 tr = make_float4(0.f, 0.f, 0.f, 0.f);
    if (in_range(rays, qx, qy)) {
      uint2 r = ld64(rays, qx, qy);
      tr = make_float4(unorm16_to_float(r.x & 0xFFFFu), unorm16_to_float(r.x >> 16), unorm16_to_float(r.y & 0xFFFFu), unorm16_to_float(r.y >> 16));
    }
    const float pu = texel_uv(qx, dim.w, dim.rw), pv = texel_uv(qy, dim.h, dim.rh);                // :117 (no +0.5)
    const float pixel_depth = fetch_d24(depth1, qx, qy);
    const f3 view_vec = reconstruct_fast(pu, pv, pixel_depth, cam);
    const float2 ne = sample_rg16unorm_fast(normal_tex, make_tapw(pu, pv, normal_tex.w, normal_tex.h));
    const f3 Nn = mul_m3f(k.normal_mat, decode_normal_fast(ne.x, ne.y));
    const f3 hit_vec = reconstruct_fast(tr.x, tr.y, tr.z, cam);
    f3 rad = mk3(0.f, 0.f, 0.f);
    if (tr.w != 1.0f) {                                                                           // is_valid_ray, :93-95,:128
      const TapW w = make_tapw(tr.x, tr.y, albedo_tex.w, albedo_tex.h);
      const uint32_t a00 = ld32(albedo_tex, w.x0, w.y0), a10 = ld32(albedo_tex, w.x1, w.y0), a01 = ld32(albedo_tex, w.x0, w.y1), a11 = ld32(albedo_tex, w.x1, w.y1);
      rad = mk3(wsum(CH(a00, 0), CH(a10, 0), CH(a01, 0), CH(a11, 0), w), wsum(CH(a00, 8), CH(a10, 8), CH(a01, 8), CH(a11, 8), w),
                wsum(CH(a00, 16), CH(a10, 16), CH(a01, 16), CH(a11, 16), w));
    }
    const f3 V = norm3f(mk3(-view_vec.x, -view_vec.y, -view_vec.z));
    const f3 L = norm3f(sub3f(hit_vec, view_vec));
    const f3 H = norm3f(mk3(V.x + L.x, V.y + L.y, V.z + L.z));
    const float fc = fminf(fmaxf(1.0f - fmaxf(dot3f(H, V), 0.0f), 0.0f), 1.0f);
    const float fc2 = fc * fc;
    const float NdotL = fmaxf(dot3f(Nn, L), 0.f), NdotV = fmaxf(dot3f(Nn, V), 0.f);
    const float l2 = NdotL * NdotL, v2 = NdotV * NdotV;
    s_a[t] = make_float4(fc2 * fc2 * fc, div_fast(1.f - l2, l2), div_fast(1.f - v2, v2), NdotV);
    s_b[t] = make_float4(rad.x, rad.y, rad.z, pixel_depth);
  }
  __syncthreads();

  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
  const int px = blockIdx.x * FT_W + tx, py = VKR_BLOCK_Y * FT_H + ty;
  if (px >= out.w || py >= out.h) return;
  const float su = texel_uv(px, dim.w, dim.rw), sv = texel_uv(py, dim.h, dim.rh);                // :39

  float roughness, metallic;
  float F0x, F0y, F0z;
  {
    const TapW t = make_tapw(su, sv, albedo_tex.w, albedo_tex.h);
    uint32_t a00 = ld32(albedo_tex, t.x0, t.y0), a10 = ld32(albedo_tex, t.x1, t.y0), a01 = ld32(albedo_tex, t.x0, t.y1), a11 = ld32(albedo_tex, t.x1, t.y1);
    uint32_t m00 = ld32(material_tex, t.x0, t.y0), m10 = ld32(material_tex, t.x1, t.y0), m01 = ld32(material_tex, t.x0, t.y1), m11 = ld32(material_tex, t.x1, t.y1);
    float ax = wsum(CH(a00, 0), CH(a10, 0), CH(a01, 0), CH(a11, 0), t);
    float ay = wsum(CH(a00, 8), CH(a10, 8), CH(a01, 8), CH(a11, 8), t);
    float az = wsum(CH(a00, 16), CH(a10, 16), CH(a01, 16), CH(a11, 16), t);
    roughness = wsum(CH(m00, 8), CH(m10, 8), CH(m01, 8), CH(m11, 8), t);
    metallic = wsum(CH(m00, 16), CH(m10, 16), CH(m01, 16), CH(m11, 16), t);
    F0x = fmaf(metallic, ax - 0.04f, 0.04f); F0y = fmaf(metallic, ay - 0.04f, 0.04f); F0z = fmaf(metallic, az - 0.04f, 0.04f);   // F0_approximation
  }
#undef CH
  const float alpha2 = roughness * roughness;
  // brdfG1(NdotV, alpha2) with the reference's swapped arguments (App. A.6 #4): the "NdotV" slot holds alpha2
  const float a4 = alpha2 * alpha2;
  const float tg_c = div_fast(1.f - a4, a4);
  const int c = (ty + 1) * (FT_W + 2) + tx + 1;
  const float center_depth = s_b[c].w;                                                          // :75
  const float bk = 1000.f * rcp_fast(center_depth);
  const bool bilateral = (k.render_flags & VKR_SSR_BILATERAL_FILTER) != 0;
  const int taps = (k.render_flags & VKR_SSR_NORMALIZE_REFLECTIONS) ? 5 : 1;
  const int offs[5] = {0, -1, FT_W + 2, 1, -(FT_W + 2)};                                         // :63-69: (0,0) (-1,0) (0,1) (1,0) (0,-1)

  float csx = 0.f, csy = 0.f, csz = 0.f, wsx = 0.f, wsy = 0.f, wsz = 0.f;
  for (int i = 0; i < taps; i++) {
    const float4 A = s_a[c + offs[i]], B = s_b[c + offs[i]];
    // ray_weight (:97-108): F * G2 / G1
    const float L1 = sqrt_fast(fmaf(alpha2, A.y, 1.f)), L2 = sqrt_fast(fmaf(alpha2, A.z, 1.f));
    const float g = (1.f + sqrt_fast(fmaf(A.w, tg_c, 1.f))) * rcp_fast(L1 + L2);                // (2/(L1+L2)) / (2/(1+sqrt(..)))
    float bw = 1.0f;
    if (bilateral) bw = fmaxf(1.f - fabsf(center_depth - B.w) * bk, 0.f);
    const float gb = g * bw;
    const float wx = fmaf(1.f - F0x, A.x, F0x) * gb, wy = fmaf(1.f - F0y, A.x, F0y) * gb, wz = fmaf(1.f - F0z, A.x, F0z) * gb;
    csx = fmaf(wx, B.x, csx); csy = fmaf(wy, B.y, csy); csz = fmaf(wz, B.z, csz);
    wsx += wx; wsy += wy; wsz += wz;
  }
  if (gmax(wsx, gmax(wsy, wsz)) < 0.001f) { wsx = 1.f; wsy = 1.f; wsz = 1.f; }
  *((uint32_t*)(out.p + (size_t)py * out.pitch) + px) = pack_rgba8unorm(csx * rcp_fast(wsx), csy * rcp_fast(wsy), csz * rcp_fast(wsz), 0.f);
}

}  // namespace vkr
using namespace vkr;

static FilterK make_filter_k(const VkrTraceParams* p) {
  FilterK k;
  k.normal_mat = make_mat(p->normal_mat);
  k.cam = make_cam(p->fovy, p->aspect, p->znear, p->zfar);
  k.render_flags = p->render_flags;
  return k;
}

extern "C" int vkr_sssr_filter_ex(const VkrTraceParams* p, const VkrImage* rays, const VkrImage* depth, const VkrImage* albedo,
                                  const VkrImage* normal, const VkrImage* material, const VkrImage* out_reflections, const VkrLaunchExt* ext, void* stream) {
  if (!p || !image_ok(rays, VKR_FORMAT_R16G16B16A16_UNORM) || !image_ok(depth, VKR_FORMAT_D24_UNORM_S8_UINT) || depth->mip_count < 2 ||
      !image_ok(albedo, VKR_FORMAT_R8G8B8A8_SRGB) || !image_ok(normal, VKR_FORMAT_R16G16_UNORM) || !image_ok(material, VKR_FORMAT_R8G8B8A8_SRGB) ||
      !image_ok(out_reflections, VKR_FORMAT_R8G8B8A8_UNORM)) {
    set_error("sssr_filter: null argument or unexpected format");
    return VKR_ERR_INVALID_ARGUMENT;
  }
  if (rays->width != out_reflections->width || rays->height != out_reflections->height) {
    set_error("sssr_filter: rays and reflections sizes differ");
    return VKR_ERR_SIZE_MISMATCH;
  }
  ExtInfo x;
  int rc = resolve_ext(ext, out_reflections->height, FT_H, &x, "sssr_filter");
  if (rc) return rc;
  const float* srgb = srgb_decode_table_device();
  if (!srgb) { set_error("sssr_filter: sRGB table upload failed"); return VKR_ERR_CUDA; }
  cudaStream_t st = (cudaStream_t)stream;
  FilterK k = make_filter_k(p);
  const FilterK* kv = nullptr;
  if (x.layers > 1) {
    if (!x.prepared()) {
      std::vector<FilterK> ks(x.layers);
      for (int i = 0; i < x.layers; i++) ks[i] = make_filter_k(p + i);
      if ((rc = upload_params(ks.data(), sizeof(FilterK), x.layers, x.scratch(), st, "sssr_filter"))) return rc;
    }
    kv = (const FilterK*)x.scratch() + x.first();
    if (x.prepare_only()) return VKR_OK;
  }
  dim3 grid((out_reflections->width + FT_W - 1) / FT_W, x.yblocks, x.layers);
  auto kern = kv ? sssr_filter_kernel<true> : sssr_filter_kernel<false>;
  kern<<<grid, 256, 0, st>>>(k, kv, LaunchX{x.yb0}, make_dim(out_reflections->width, out_reflections->height), make_mip(rays, 0, x, 0), make_mip(depth, 1, x, 1),
                             make_mip(albedo, 0, x, 2), make_mip(normal, 0, x, 3), make_mip(material, 0, x, 4), make_mip(out_reflections, 0, x, 5), srgb);
  return check_launch("sssr_filter_kernel");
}

extern "C" int vkr_sssr_filter(const VkrTraceParams* p, const VkrImage* rays, const VkrImage* depth, const VkrImage* albedo,
                               const VkrImage* normal, const VkrImage* material, const VkrImage* out_reflections, void* stream) {
  return vkr_sssr_filter_ex(p, rays, depth, albedo, normal, material, out_reflections, nullptr, stream);
}
