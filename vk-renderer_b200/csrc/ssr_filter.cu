// ssr_filter.cu — 5-tap BRDF-weighted resolve of the traced rays.
// Replaces task "SSSR_filter" (reference src/advanced_ssr.cpp:308-369,
// shaders/advanced_ssr/filter.comp:36-149).  Quirks kept on purpose:
// uv = pixel / size without +0.5 (#5) and the swapped brdfG1 arguments (#4).
#include "brdf.cuh"

namespace vkr {

struct FilterK {
  Mat4 normal_mat;
  CamK cam;
  uint32_t render_flags;
};

__global__ void __launch_bounds__(256) sssr_filter_kernel(FilterK k, Mip rays, Mip depth1, Mip albedo_tex, Mip normal_tex, Mip material_tex, Mip out,
                                                          const float* __restrict__ srgb) {
  const int px = blockIdx.x * 32 + (threadIdx.x & 31), py = blockIdx.y * 8 + (threadIdx.x >> 5);
  if (px >= out.w || py >= out.h) return;
  const float tsx = (float)out.w, tsy = (float)out.h;
  const float su = (float)px / tsx, sv = (float)py / tsy;                                     // :39

  float4 material = sample_srgb8(material_tex, su, sv, srgb);
  const float metallic = material.z, roughness = material.y;
  float4 a4 = sample_srgb8(albedo_tex, su, sv, srgb);
  f3 F0 = mix3(mk3(0.04f, 0.04f, 0.04f), mk3(a4.x, a4.y, a4.z), metallic);
  f3 color_sum = mk3(0, 0, 0), weight_sum = mk3(0, 0, 0);
  const float center_depth = fetch_d24(depth1, px, py);                                       // :75
  const float alpha2 = FM(roughness, roughness);
  const int taps = (k.render_flags & VKR_SSR_NORMALIZE_REFLECTIONS) ? 5 : 1;
  const int offx[5] = {0, -1, 0, 1, 0}, offy[5] = {0, 0, 1, 0, -1};                            // :63-69

  for (int i = 0; i < taps; i++) {                                                              // process_pixel, :110-149
    const int qx = px + offx[i], qy = py + offy[i];
    float4 tr = make_float4(0.f, 0.f, 0.f, 0.f);
    if (in_range(rays, qx, qy)) {
      uint2 r = ld64(rays, qx, qy);
      tr = make_float4(unorm16_to_float(r.x & 0xFFFFu), unorm16_to_float(r.x >> 16), unorm16_to_float(r.y & 0xFFFFu), unorm16_to_float(r.y >> 16));
    }
    const float pu = (float)qx / tsx, pv = (float)qy / tsy;
    const float pixel_depth = fetch_d24(depth1, qx, qy);
    f3 view_vec = reconstruct_view_vec(pu, pv, pixel_depth, k.cam);
    f3 Nn = mul_m4(k.normal_mat, sample_gbuffer_normal(normal_tex, pu, pv), 0.f);
    f3 hit_vec = reconstruct_view_vec(tr.x, tr.y, tr.z, k.cam);
    f3 radiance = mk3(0, 0, 0);
    if (tr.w != 1.0f) { float4 c = sample_srgb8(albedo_tex, tr.x, tr.y, srgb); radiance = mk3(c.x, c.y, c.z); }
    f3 V = -norm3(view_vec);
    f3 L = norm3(hit_vec - view_vec);
    // ray_weight, :97-108
    f3 H = norm3(V + L);
    float fs = powf(gclamp(FS(1.0f, gmax(dot3(H, V), 0.0f)), 0.0f, 1.0f), 5.0f);
    f3 one = mk3(1.f, 1.f, 1.f);
    f3 F = F0 + (one - F0) * fs;
    float NdotL = gmax(dot3(Nn, L), 0.f);
    float NdotV = gmax(dot3(Nn, V), 0.f);
    float G2 = brdfG2(NdotL, NdotV, alpha2);
    float G1 = brdfG1(NdotV, alpha2);      // argument order as written in the reference
    f3 weight = (F * G2) / G1;
    float bilateral_weight = 1.0f;
    if (k.render_flags & VKR_SSR_BILATERAL_FILTER) bilateral_weight = gmax(FS(1.f, FM(1000.f, fabsf(FS(center_depth, pixel_depth))) / center_depth), 0.f);
    weight = weight * bilateral_weight;
    color_sum = color_sum + weight * radiance;
    weight_sum = weight_sum + weight;
  }
  if (gmax(weight_sum.x, gmax(weight_sum.y, weight_sum.z)) < 0.001f) weight_sum = mk3(1, 1, 1);
  *((uint32_t*)(out.p + (size_t)py * out.pitch) + px) =
      pack_rgba8unorm(color_sum.x / weight_sum.x, color_sum.y / weight_sum.y, color_sum.z / weight_sum.z, 0.f);
}

}  // namespace vkr
using namespace vkr;

extern "C" int vkr_sssr_filter(const VkrTraceParams* p, const VkrImage* rays, const VkrImage* depth, const VkrImage* albedo,
                               const VkrImage* normal, const VkrImage* material, const VkrImage* out_reflections, void* stream) {
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
  FilterK k;
  k.normal_mat = make_mat(p->normal_mat);
  k.cam = make_cam(p->fovy, p->aspect, p->znear, p->zfar);
  k.render_flags = p->render_flags;
  const float* srgb = srgb_decode_table_device();
  if (!srgb) { set_error("sssr_filter: sRGB table upload failed"); return VKR_ERR_CUDA; }
  dim3 grid((out_reflections->width + 31) / 32, (out_reflections->height + 7) / 8);
  sssr_filter_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(k, make_mip(rays, 0), make_mip(depth, 1), make_mip(albedo, 0), make_mip(normal, 0),
                                                             make_mip(material, 0), make_mip(out_reflections, 0), srgb);
  return check_launch("sssr_filter_kernel");
}
