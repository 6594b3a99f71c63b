// shading.cu — deferred PBR point-light shading + depth-aware 2x2 pick of half-res AO and
// reflections + split-sum BRDF LUT.  Replaces task "DeferedShading"
// (reference src/defered_shading.cpp:47-118, shaders/defered_shading/shader.frag:41-130).
// Output is R8G8B8A8_SRGB (linear -> sRGB on store, reference src/main.cpp:289-291).
#include "brdf.cuh"

namespace vkr {

struct ShadeK {
  Mat4 inverse_camera;
  CamK cam;
  float min_r, max_r;
  uint32_t show_ao;
};

// brdf.glsl:6-13,:31-38,:49-56
__device__ __forceinline__ float pow5(float x) { return powf(x, 5.0f); }
__global__ void __launch_bounds__(256) deferred_shading_kernel(ShadeK k, Mip albedo_tex, Mip normal_tex, Mip material_tex, Mip depth0, Mip depth1,
                                                               Mip occlusion_tex, Mip brdf_tex, Mip reflections_tex, Mip out,
                                                               const float* __restrict__ srgb) {
  const int px = blockIdx.x * 32 + (threadIdx.x & 31), py = blockIdx.y * 8 + (threadIdx.x >> 5);
  if (px >= out.w || py >= out.h) return;
  const float su = ((float)px + 0.5f) / (float)out.w, sv = ((float)py + 0.5f) / (float)out.h;

  f3 N = sample_gbuffer_normal(normal_tex, su, sv);
  float4 a4 = sample_srgb8(albedo_tex, su, sv, srgb);
  f3 albedo = mk3(a4.x, a4.y, a4.z);
  float4 material = sample_srgb8(material_tex, su, sv, srgb);
  float depth = sample_d24(depth0, su, sv);

  // sample_ocllusion_ssr, shader.frag:103-130: of the 2x2 half-res neighbourhood pick the
  // texel whose mip-1 depth is closest, then sample AO / reflections with the same offset.
  float dl0 = fabsf(sample_d24(depth1, su, sv, 0, 0) - depth), dl1 = fabsf(sample_d24(depth1, su, sv, 1, 0) - depth);
  float dl2 = fabsf(sample_d24(depth1, su, sv, 0, 1) - depth), dl3 = fabsf(sample_d24(depth1, su, sv, 1, 1) - depth);
  float min_delta = gmin(gmin(dl0, dl1), gmin(dl2, dl3));
  int ox, oy;
  if (min_delta == dl0) { ox = 0; oy = 0; }
  else if (min_delta == dl1) { ox = 1; oy = 0; }
  else if (min_delta == dl2) { ox = 0; oy = 1; }
  else { ox = 1; oy = 1; }
  float occlusion = sample_rg16f(occlusion_tex, su, sv, ox, oy).x;
  f3 reflection;
  {
    Tap t = make_tap(su, sv, reflections_tex.w, reflections_tex.h, ox, oy);
    f3 a = unpack_rgb8unorm(ld32(reflections_tex, t.x0, t.y0)), b = unpack_rgb8unorm(ld32(reflections_tex, t.x1, t.y0));
    f3 c = unpack_rgb8unorm(ld32(reflections_tex, t.x0, t.y1)), d = unpack_rgb8unorm(ld32(reflections_tex, t.x1, t.y1));
    reflection = mk3(bilerp(a.x, b.x, c.x, d.x, t.wx, t.wy), bilerp(a.y, b.y, c.y, d.y, t.wx, t.wy), bilerp(a.z, b.z, c.z, d.z, t.wx, t.wy));
  }

  f3 camera_view_vec = reconstruct_view_vec(su, sv, depth, k.cam);
  f3 world_pos = mul_m4(k.inverse_camera, camera_view_vec, 1.f);
  f3 camera_pos = mk3(k.inverse_camera.m[12], k.inverse_camera.m[13], k.inverse_camera.m[14]);

  const float metallic = gmix(0.1f, 1.0f, material.z);
  const float roughness = material.y;
  const f3 V = norm3(camera_pos - world_pos);
  f3 F0 = mix3(mk3(0.04f, 0.04f, 0.04f), albedo, metallic);

  const f3 LIGHT_POS = mk3(-1.85867f, 5.81832f, -0.247114f);
  f3 Lv = LIGHT_POS - world_pos;
  f3 L = norm3(Lv);
  f3 H = norm3(V + L);
  float light_distance = len3(Lv);
  float rad = FM(0.1f, gmin(100 / FM(light_distance, light_distance), 100.0f));

  float NdotL = gmax(dot3(N, L), 0.f);
  float NdotV = gmax(dot3(N, V), 0.f);
  float NDF = DistributionGGX(N, H, roughness);
  float G = brdfG2(NdotV, NdotL, FM(roughness, roughness));
  float fs = pow5(gclamp(FS(1.0f, gmax(dot3(H, V), 0.0f)), 0.0f, 1.0f));
  f3 one = mk3(1.f, 1.f, 1.f);
  f3 F = F0 + (one - F0) * fs;
  f3 kD = (one - F) * FS(1.f, metallic);
  f3 specular = (FM(NDF, G) * F) / FA(FM(FM(4.0f, NdotV), NdotL), 0.0001f);
  float biased_rougness = gmix(k.min_r, k.max_r, roughness);
  float2 ssr_brdf = sample_rg16f(brdf_tex, biased_rougness, NdotV);

  f3 Lo = ((kD * albedo) / VKR_PI + specular) * rad * NdotL;
  Lo = Lo + reflection * (F0 * ssr_brdf.x + mk3(ssr_brdf.y, ssr_brdf.y, ssr_brdf.y));
  f3 color = occlusion * (0.6f * albedo + Lo);
  if (k.show_ao != 0) color = mk3(occlusion, occlusion, occlusion);
  *((uint32_t*)(out.p + (size_t)py * out.pitch) + px) = pack_rgba8unorm(srgb_encode(color.x), srgb_encode(color.y), srgb_encode(color.z), 0.f);
}

}  // namespace vkr
using namespace vkr;

extern "C" int vkr_deferred_shading(const VkrShadingParams* p, const VkrImage* albedo, const VkrImage* normal, const VkrImage* material,
                                    const VkrImage* depth, const VkrImage* occlusion, const VkrImage* brdf_lut, const VkrImage* reflections,
                                    const VkrImage* out_color, void* stream) {
  if (!p || !image_ok(albedo, VKR_FORMAT_R8G8B8A8_SRGB) || !image_ok(normal, VKR_FORMAT_R16G16_UNORM) || !image_ok(material, VKR_FORMAT_R8G8B8A8_SRGB) ||
      !image_ok(depth, VKR_FORMAT_D24_UNORM_S8_UINT) || depth->mip_count < 2 || !image_ok(occlusion, VKR_FORMAT_R16G16_SFLOAT) ||
      !image_ok(brdf_lut, VKR_FORMAT_R16G16_SFLOAT) || !image_ok(reflections, VKR_FORMAT_R8G8B8A8_UNORM) || !image_ok(out_color, VKR_FORMAT_R8G8B8A8_SRGB)) {
    set_error("deferred_shading: null argument or unexpected format");
    return VKR_ERR_INVALID_ARGUMENT;
  }
  ShadeK k;
  k.inverse_camera = make_mat(p->inverse_camera);
  k.cam = make_cam(p->fovy, p->aspect, p->znear, p->zfar);
  k.min_r = p->min_max_roughness[0]; k.max_r = p->min_max_roughness[1]; k.show_ao = p->show_ao;
  const float* srgb = srgb_decode_table_device();
  if (!srgb) { set_error("deferred_shading: sRGB table upload failed"); return VKR_ERR_CUDA; }
  dim3 grid((out_color->width + 31) / 32, (out_color->height + 7) / 8);
  deferred_shading_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(k, make_mip(albedo, 0), make_mip(normal, 0), make_mip(material, 0), make_mip(depth, 0),
                                                                  make_mip(depth, 1), make_mip(occlusion, 0), make_mip(brdf_lut, 0),
                                                                  make_mip(reflections, 0), make_mip(out_color, 0), srgb);
  return check_launch("deferred_shading_kernel");
}
