// shading.cu — deferred PBR point-light shading + depth-aware 2x2 pick of half-res AO and
// reflections + split-sum BRDF LUT.  Replaces task "DeferedShading"
// (reference src/defered_shading.cpp:47-118, shaders/defered_shading/shader.frag:41-130).
// Output is R8G8B8A8_SRGB (linear -> sRGB on store, reference src/main.cpp:289-291).
#include <vector>
#include "brdf.cuh"

#ifndef SHADE_MIN_CTAS
#define SHADE_MIN_CTAS 8   // resident CTAs per SM the register allocation must allow (tools/variant.sh sweeps it)
#endif
namespace vkr {

struct ShadeK {
  Mat4 inverse_camera;
  CamK cam;
  float min_r, max_r;
  uint32_t show_ao;
};

// Colour-only arithmetic (no discrete decision depends on it): FMA contraction, reciprocal-multiply and
// the MUFU-based pow are allowed here; everything that SELECTS something (texel footprints, the
// closest-depth pick) stays on the unfused helpers of common.cuh.
struct c3 { float x, y, z; };
__device__ __forceinline__ c3 C3(float x, float y, float z) { c3 r; r.x = x; r.y = y; r.z = z; return r; }
__device__ __forceinline__ c3 C3(f3 v) { return C3(v.x, v.y, v.z); }
__device__ __forceinline__ c3 operator+(c3 a, c3 b) { return C3(a.x + b.x, a.y + b.y, a.z + b.z); }
__device__ __forceinline__ c3 operator-(c3 a, c3 b) { return C3(a.x - b.x, a.y - b.y, a.z - b.z); }
__device__ __forceinline__ c3 operator*(c3 a, c3 b) { return C3(a.x * b.x, a.y * b.y, a.z * b.z); }
__device__ __forceinline__ c3 operator*(c3 a, float s) { return C3(a.x * s, a.y * s, a.z * s); }
__device__ __forceinline__ c3 operator*(float s, c3 a) { return C3(a.x * s, a.y * s, a.z * s); }

__device__ __forceinline__ float srgb_encode_fast(float l) {       // linear -> sRGB OETF, MUFU lg2/ex2
  if (!(l > 0.0f)) return 0.0f;
  if (l >= 1.0f) return 1.0f;
  if (l < 0.0031308f) return 12.92f * l;
  return 1.055f * exp2f(__log2f(l) * (1.0f / 2.4f)) - 0.055f;
}

// SAME: albedo / normal / material have the output's extent, so their bilinear footprint at the pixel's own
// centre is the texel itself (common.cuh tap_is_texel).
// HALF: depth mip 1, the AO image and the reflections all have the extent (W/2, H/2) of the reference (W, H even).  The
// 32x8 pixel tile then touches a 19x7 block of half-res texels; each is fetched and decoded once per CTA into shared memory
// ({reflection rgb, AO} and the exact mip-1 depth) instead of ~17 times by the threads whose 3x3 / 2x2 footprints cover it.
#define SH_HW 19
#define SH_HH 7
template <bool SAME, bool LAYERED, bool HALF>
__global__ void __launch_bounds__(256, SHADE_MIN_CTAS) deferred_shading_kernel(ShadeK k0, const ShadeK* __restrict__ kv, LaunchX lx, Dim dim, Mip albedo_tex, Mip normal_tex,
                                                               Mip material_tex, Mip depth0, Mip depth1, Mip occlusion_tex, Mip brdf_tex, Mip reflections_tex,
                                                               Mip out, const float* __restrict__ srgb) {
  ShadeK kl;
  if (LAYERED) {
    const int z = blockIdx.z;
    kl = kv[z];
    albedo_tex = layer_of(albedo_tex, z); normal_tex = layer_of(normal_tex, z); material_tex = layer_of(material_tex, z);
    depth0 = layer_of(depth0, z); depth1 = layer_of(depth1, z); occlusion_tex = layer_of(occlusion_tex, z);
    brdf_tex = layer_of(brdf_tex, z); reflections_tex = layer_of(reflections_tex, z); out = layer_of(out, z);
  }
  const ShadeK& k = LAYERED ? kl : k0;
  const int px = blockIdx.x * 32 + (threadIdx.x & 31), py = VKR_BLOCK_Y * 8 + (threadIdx.x >> 5);
  __shared__ float4 s_ra[HALF ? SH_HH * SH_HW : 1];
  __shared__ float s_d1[HALF ? SH_HH * SH_HW : 1];
  const int hx0 = (int)blockIdx.x * 16 - 1, hy0 = VKR_BLOCK_Y * 4 - 1;      // half-res origin of the staged block
  if (HALF) {
    if (threadIdx.x < SH_HH * SH_HW) {
      const int ly = threadIdx.x / SH_HW, lx = threadIdx.x - ly * SH_HW;
      const int qx = min(max(hx0 + lx, 0), depth1.w - 1), qy = min(max(hy0 + ly, 0), depth1.h - 1);   // clamp-to-edge per index
      const f3 c = unpack_rgb8unorm(ld32(reflections_tex, qx, qy));
      s_ra[threadIdx.x] = make_float4(c.x, c.y, c.z, unpack_rg16f(ld32(occlusion_tex, qx, qy)).x);
      s_d1[threadIdx.x] = d24_to_float(ld32(depth1, qx, qy));
    }
    __syncthreads();
  }
  if (px >= out.w || py >= out.h) return;
  const float su = pixel_uv(px, dim.w, dim.rw), sv = pixel_uv(py, dim.h, dim.rh);
#define CH(k, s) __ldg(srgb + (((k) >> (s)) & 255u))

  f3 N;
  c3 albedo;
  float roughness, mat_b;
  if (SAME) {
    float2 e = unpack_rg16unorm(ld32(normal_tex, px, py));
    N = decode_normal_fast(e.x, e.y);
    const uint32_t a = ld32(albedo_tex, px, py), m = ld32(material_tex, px, py);
    albedo = C3(CH(a, 0), CH(a, 8), CH(a, 16));
    roughness = CH(m, 8); mat_b = CH(m, 16);
    if (((m >> 8) & 255u) == 0u) {
      // roughness == 0 is a singular point of the BRDF (alpha2 * x / 0 is NaN for alpha2 == 0 but +inf for any
      // alpha2 > 0): a sky / cleared texel beside geometry gets a ~1e-5-weight bleed from its neighbours in the
      // reference's bilinear fetch, which decides between the two.  Evaluate the real footprint for these texels.
      const Tap t = make_tap(su, sv, material_tex.w, material_tex.h);
      roughness = bilerp(CH(ld32(material_tex, t.x0, t.y0), 8), CH(ld32(material_tex, t.x1, t.y0), 8), CH(ld32(material_tex, t.x0, t.y1), 8),
                         CH(ld32(material_tex, t.x1, t.y1), 8), t.wx, t.wy);
    }
  } else {
    N = sample_gbuffer_normal_fast(normal_tex, su, sv);
    // albedo (rgb), material (.g roughness, .b metallic) share one footprint (same size, same uv)
    const TapW t = make_tapw(su, sv, albedo_tex.w, albedo_tex.h);
    uint32_t a00 = ld32(albedo_tex, t.x0, t.y0), a10 = ld32(albedo_tex, t.x1, t.y0), a01 = ld32(albedo_tex, t.x0, t.y1), a11 = ld32(albedo_tex, t.x1, t.y1);
    uint32_t m00 = ld32(material_tex, t.x0, t.y0), m10 = ld32(material_tex, t.x1, t.y0), m01 = ld32(material_tex, t.x0, t.y1), m11 = ld32(material_tex, t.x1, t.y1);
    albedo.x = wsum(CH(a00, 0), CH(a10, 0), CH(a01, 0), CH(a11, 0), t);
    albedo.y = wsum(CH(a00, 8), CH(a10, 8), CH(a01, 8), CH(a11, 8), t);
    albedo.z = wsum(CH(a00, 16), CH(a10, 16), CH(a01, 16), CH(a11, 16), t);
    roughness = wsum(CH(m00, 8), CH(m10, 8), CH(m01, 8), CH(m11, 8), t);
    mat_b = wsum(CH(m00, 16), CH(m10, 16), CH(m01, 16), CH(m11, 16), t);
  }
#undef CH
  // Exact (unfused) from here to the pick: on planar surfaces the four deltas are near-ties and the argmin
  // decides which half-res texel feeds AO / reflections.
  const float depth = sample_d24(depth0, su, sv);

  // sample_ocllusion_ssr, shader.frag:103-130.  The four textureLodOffset(depth, uv, 1, off) footprints
  // cover a 3x3 block of half-res texels: fetch and decode each once.
  int ox, oy, hbase = 0;
  Tap t1 = make_tap(su, sv, depth1.w, depth1.h);
  {
    float x = __fsub_rn(__fmul_rn(su, (float)depth1.w), 0.5f), y = __fsub_rn(__fmul_rn(sv, (float)depth1.h), 0.5f);
    const int ix = (int)floorf(x), iy = (int)floorf(y);
    float d[3][3];
    if (HALF) {
      hbase = (iy - hy0) * SH_HW + (ix - hx0);
#pragma unroll
      for (int j = 0; j < 3; j++)
#pragma unroll
        for (int i = 0; i < 3; i++) d[j][i] = s_d1[hbase + j * SH_HW + i];
    } else {
      int cx[3], cy[3];
#pragma unroll
      for (int i = 0; i < 3; i++) { cx[i] = min(max(ix + i, 0), depth1.w - 1); cy[i] = min(max(iy + i, 0), depth1.h - 1); }
#pragma unroll
      for (int j = 0; j < 3; j++)
#pragma unroll
        for (int i = 0; i < 3; i++) d[j][i] = d24_to_float(ld32(depth1, cx[i], cy[j]));
    }
    // the four bilerp(...) of the reference on packed fp32: row j gives the horizontal mixes of its two texel pairs at
    // once, then rows (0,1) and (1,2) are mixed vertically — each lane rounds like the scalar gmix chain
    const f2 one = one2();
    const float omx = FS(1.0f, t1.wx), omy = FS(1.0f, t1.wy);
    const f2 h0 = gmix2(pk(d[0][0], d[0][1]), pk(d[0][1], d[0][2]), omx, t1.wx, one);
    const f2 h1 = gmix2(pk(d[1][0], d[1][1]), pk(d[1][1], d[1][2]), omx, t1.wx, one);
    const f2 h2 = gmix2(pk(d[2][0], d[2][1]), pk(d[2][1], d[2][2]), omx, t1.wx, one);
    float dl0, dl1, dl2, dl3;
    unpk(sub2(gmix2(h0, h1, omy, t1.wy, one), dup(depth)), dl0, dl1);
    unpk(sub2(gmix2(h1, h2, omy, t1.wy, one), dup(depth)), dl2, dl3);
    dl0 = fabsf(dl0); dl1 = fabsf(dl1); dl2 = fabsf(dl2); dl3 = fabsf(dl3);
    float min_delta = gmin(gmin(dl0, dl1), gmin(dl2, dl3));
    if (min_delta == dl0) { ox = 0; oy = 0; }
    else if (min_delta == dl1) { ox = 1; oy = 0; }
    else if (min_delta == dl2) { ox = 0; oy = 1; }
    else { ox = 1; oy = 1; }
  }
  // --- value-only from here on ---
  float occlusion;
  c3 reflection;
  if (HALF) {
    const TapW t = make_tapw(t1);                                          // same extent: same weights; indices come from the staged block
    const int b = hbase + oy * SH_HW + ox;
    const float4 a00 = s_ra[b], a10 = s_ra[b + 1], a01 = s_ra[b + SH_HW], a11 = s_ra[b + SH_HW + 1];
    occlusion = wsum(a00.w, a10.w, a01.w, a11.w, t);
    reflection = C3(wsum(a00.x, a10.x, a01.x, a11.x, t), wsum(a00.y, a10.y, a01.y, a11.y, t), wsum(a00.z, a10.z, a01.z, a11.z, t));
  } else {
    // occlusion and reflections have the extent of depth mip 1 in the reference (W/2 x H/2): one footprint serves both
    const bool shared_fp = occlusion_tex.w == reflections_tex.w && occlusion_tex.h == reflections_tex.h;
    const TapW t = make_tapw(su, sv, occlusion_tex.w, occlusion_tex.h, ox, oy);
    occlusion = wsum(unpack_rg16f(ld32(occlusion_tex, t.x0, t.y0)).x, unpack_rg16f(ld32(occlusion_tex, t.x1, t.y0)).x,
                     unpack_rg16f(ld32(occlusion_tex, t.x0, t.y1)).x, unpack_rg16f(ld32(occlusion_tex, t.x1, t.y1)).x, t);
    const TapW tr = shared_fp ? t : make_tapw(su, sv, reflections_tex.w, reflections_tex.h, ox, oy);
    f3 a = unpack_rgb8unorm(ld32(reflections_tex, tr.x0, tr.y0)), b = unpack_rgb8unorm(ld32(reflections_tex, tr.x1, tr.y0));
    f3 c = unpack_rgb8unorm(ld32(reflections_tex, tr.x0, tr.y1)), d = unpack_rgb8unorm(ld32(reflections_tex, tr.x1, tr.y1));
    reflection = C3(wsum(a.x, b.x, c.x, d.x, tr), wsum(a.y, b.y, c.y, d.y, tr), wsum(a.z, b.z, c.z, d.z, tr));
  }

  const CamF cam = make_camf(k.cam);
  const f3 world_pos = mul_m4f(k.inverse_camera, reconstruct_fast(su, sv, depth, cam));
  const f3 camera_pos = mk3(k.inverse_camera.m[12], k.inverse_camera.m[13], k.inverse_camera.m[14]);

  const float metallic = fmaf(0.9f, mat_b, 0.1f);                           // mix(0.1, 1.0, material.b)
  const f3 V = norm3f(sub3f(camera_pos, world_pos));
  const f3 LIGHT_POS = mk3(-1.85867f, 5.81832f, -0.247114f);                 // shader.frag:37
  const f3 Lv = sub3f(LIGHT_POS, world_pos);
  const float ld2 = dot3f(Lv, Lv);                                          // light_distance^2
  const float ild = rsqrtf(ld2);
  const f3 L = mk3(Lv.x * ild, Lv.y * ild, Lv.z * ild);
  const f3 H = norm3f(mk3(V.x + L.x, V.y + L.y, V.z + L.z));
  const float rad = 0.1f * fminf(div_fast(100.f, ld2), 100.0f);          // LIGHT_RADIANCE * min(100/d^2, 100)

  const float NdotL = fmaxf(dot3f(N, L), 0.f), NdotV = fmaxf(dot3f(N, V), 0.f);
  const c3 one = C3(1.f, 1.f, 1.f);
  const c3 F0 = C3(0.04f, 0.04f, 0.04f) * (1.0f - metallic) + albedo * metallic;
  // DistributionGGX(N, H, roughness), brdf.glsl:31-38
  const float NoH = dot3f(N, H), a2 = roughness * roughness, NoH2 = NoH * NoH;
  const float den = fmaf(NoH2, a2, 1.f - NoH2);
  const float NDF = (NoH2 > 0 ? a2 : 0.f) * rcp_fast(VKR_PI * den * den);
  // brdfG2(NdotV, NdotL, roughness^2), brdf.glsl:49-56:  2 / (sqrt(1 + a(1-v^2)/v^2) + sqrt(1 + a(1-l^2)/l^2))
  const float v2 = NdotV * NdotV, l2 = NdotL * NdotL;
  const float L1 = sqrt_fast(1.f + div_fast(a2 * (1.f - v2), v2)), L2 = sqrt_fast(1.f + div_fast(a2 * (1.f - l2), l2));
  const float G = div_fast(2.0f, L1 + L2);
  float fc = fminf(fmaxf(1.0f - fmaxf(dot3f(H, V), 0.0f), 0.0f), 1.0f);
  float fc2 = fc * fc;
  const float fs = fc2 * fc2 * fc;                                          // pow(x, 5)
  const c3 F = F0 + (one - F0) * fs;
  const c3 kD = (one - F) * (1.f - metallic);
  const float inv_den = rcp_fast(4.0f * NdotV * NdotL + 0.0001f);
  const c3 specular = F * (NDF * G * inv_den);
  const float biased_rougness = fmaf(roughness, k.max_r - k.min_r, k.min_r);   // mix(min, max, roughness)
  const float2 ssr_brdf = sample_rg16f_fast(brdf_tex, make_tapw(biased_rougness, NdotV, brdf_tex.w, brdf_tex.h));

  c3 Lo = (kD * albedo * (1.0f / VKR_PI) + specular) * (rad * NdotL);
  Lo = Lo + reflection * (F0 * ssr_brdf.x + C3(ssr_brdf.y, ssr_brdf.y, ssr_brdf.y));
  c3 color = occlusion * (0.6f * albedo + Lo);
  if (k.show_ao != 0) color = C3(occlusion, occlusion, occlusion);
  *((uint32_t*)(out.p + (size_t)py * out.pitch) + px) =
      pack_rgba8unorm(srgb_encode_fast(color.x), srgb_encode_fast(color.y), srgb_encode_fast(color.z), 0.f);
}

}  // namespace vkr
using namespace vkr;

static ShadeK make_shade_k(const VkrShadingParams* p) {
  ShadeK k;
  k.inverse_camera = make_mat(p->inverse_camera);
  k.cam = make_cam(p->fovy, p->aspect, p->znear, p->zfar);
  k.min_r = p->min_max_roughness[0]; k.max_r = p->min_max_roughness[1]; k.show_ao = p->show_ao;
  return k;
}

extern "C" int vkr_deferred_shading_ex(const VkrShadingParams* p, const VkrImage* albedo, const VkrImage* normal, const VkrImage* material,
                                       const VkrImage* depth, const VkrImage* occlusion, const VkrImage* brdf_lut, const VkrImage* reflections,
                                       const VkrImage* out_color, const VkrLaunchExt* ext, void* stream) {
  if (!p || !image_ok(albedo, VKR_FORMAT_R8G8B8A8_SRGB) || !image_ok(normal, VKR_FORMAT_R16G16_UNORM) || !image_ok(material, VKR_FORMAT_R8G8B8A8_SRGB) ||
      !image_ok(depth, VKR_FORMAT_D24_UNORM_S8_UINT) || depth->mip_count < 2 || !image_ok(occlusion, VKR_FORMAT_R16G16_SFLOAT) ||
      !image_ok(brdf_lut, VKR_FORMAT_R16G16_SFLOAT) || !image_ok(reflections, VKR_FORMAT_R8G8B8A8_UNORM) || !image_ok(out_color, VKR_FORMAT_R8G8B8A8_SRGB)) {
    set_error("deferred_shading: null argument or unexpected format");
    return VKR_ERR_INVALID_ARGUMENT;
  }
  ExtInfo x;
  int rc = resolve_ext(ext, out_color->height, 8, &x, "deferred_shading");
  if (rc) return rc;
  const float* srgb = srgb_decode_table_device();
  if (!srgb) { set_error("deferred_shading: sRGB table upload failed"); return VKR_ERR_CUDA; }
  cudaStream_t st = (cudaStream_t)stream;
  ShadeK k = make_shade_k(p);
  const ShadeK* kv = nullptr;
  if (x.layers > 1) {
    if (!x.prepared()) {
      std::vector<ShadeK> ks(x.layers);
      for (int i = 0; i < x.layers; i++) ks[i] = make_shade_k(p + i);
      if ((rc = upload_params(ks.data(), sizeof(ShadeK), x.layers, x.scratch(), st, "deferred_shading"))) return rc;
    }
    kv = (const ShadeK*)x.scratch() + x.first();
    if (x.prepare_only()) return VKR_OK;
  }
  dim3 grid((out_color->width + 31) / 32, x.yblocks, x.layers);
  const Dim dim = make_dim(out_color->width, out_color->height);
  const bool same = same_extent(albedo, out_color) && same_extent(normal, out_color) && same_extent(material, out_color) &&
                    out_color->width <= 8192 && out_color->height <= 8192;
  const uint32_t w2 = out_color->width / 2, h2 = out_color->height / 2;
  const bool half = same && (out_color->width % 2 == 0) && (out_color->height % 2 == 0) && occlusion->width == w2 && occlusion->height == h2 &&
                    reflections->width == w2 && reflections->height == h2 && mip_dim(depth->width, 1) == (int)w2 && mip_dim(depth->height, 1) == (int)h2 &&
                    same_extent(depth, out_color);
  auto kern = half ? (kv ? deferred_shading_kernel<true, true, true> : deferred_shading_kernel<true, false, true>)
              : same ? (kv ? deferred_shading_kernel<true, true, false> : deferred_shading_kernel<true, false, false>)
                     : (kv ? deferred_shading_kernel<false, true, false> : deferred_shading_kernel<false, false, false>);
  kern<<<grid, 256, 0, st>>>(k, kv, LaunchX{x.yb0}, dim, make_mip(albedo, 0, x, 0), make_mip(normal, 0, x, 1), make_mip(material, 0, x, 2),
                             make_mip(depth, 0, x, 3), make_mip(depth, 1, x, 3), make_mip(occlusion, 0, x, 4), make_mip(brdf_lut, 0, x, 5),
                             make_mip(reflections, 0, x, 6), make_mip(out_color, 0, x, 7), srgb);
  return check_launch("deferred_shading_kernel");
}

extern "C" int vkr_deferred_shading(const VkrShadingParams* p, const VkrImage* albedo, const VkrImage* normal, const VkrImage* material,
                                    const VkrImage* depth, const VkrImage* occlusion, const VkrImage* brdf_lut, const VkrImage* reflections,
                                    const VkrImage* out_color, void* stream) {
  return vkr_deferred_shading_ex(p, albedo, normal, material, depth, occlusion, brdf_lut, reflections, out_color, nullptr, stream);
}
