// alt.cu — the renderer's earlier screen-space passes, kept beside the live ones:
//   task "SSAO"  16 view-space sphere taps compared against the depth buffer       (reference src/ssao.cpp:54-96,
//                shaders/ssao/shader.frag:21-41), R8_UNORM output;
//   task "SSR"   mirror reflection, Hi-Z march of <= 100 cells from mip 0, three rejections, border fade, GGX(N,N) weight
//                (reference src/ssr.cpp:10-73, shaders/ssr/shader.frag:28-98, include/screen_trace.glsl:51-100).
// Both are fragment passes over a fullscreen triangle in the reference; here one thread per output pixel.
// SSAO's output is a count of 16 strict comparisons and the SSR march picks texels, so both are evaluated with the
// oracle's unfused arithmetic throughout (common.cuh precision policy: discontinuous choices).
#include <vector>
#include "brdf.cuh"
#include "march.cuh"

namespace vkr {

struct SsaoK {
  Mat4 projection;
  CamK cam;
};
struct SsaoSamples { float4 s[16]; };   // shared by every layer: one sample set per pass object (ssao.cpp:34-52)

template <bool LAYERED>
__global__ void __launch_bounds__(256) ssao_kernel(SsaoK k0, const SsaoK* __restrict__ kv, SsaoSamples S, LaunchX lx, Dim dim, Mip depth, Mip out) {
  SsaoK kl;
  if (LAYERED) {
    const int z = blockIdx.z;
    kl = kv[z];
    depth = layer_of(depth, z); out = layer_of(out, z);
  }
  const SsaoK& k = LAYERED ? kl : k0;
  const int px = blockIdx.x * 32 + (threadIdx.x & 31), py = VKR_BLOCK_Y * 8 + (threadIdx.x >> 5);
  if (px >= out.w || py >= out.h) return;
  const float su = pixel_uv(px, dim.w, dim.rw), sv = pixel_uv(py, dim.h, dim.rh);
  const float frag_depth = sample_d24(depth, su, sv);
  const f3 camera_pos = reconstruct_view_vec(su, sv, frag_depth, k.cam);
  float sum = 0.f;
#pragma unroll 4
  for (int i = 0; i < 16; i++) {
    const f3 pos = camera_pos + 0.05f * mk3(S.s[i].x, S.s[i].y, S.s[i].z);
    const float4 clip = mul_m4_4(k.projection, pos, 1.0f);
    const float nx = clip.x / clip.w, ny = clip.y / clip.w, nz = clip.z / clip.w;
    const float sample_depth = sample_d24(depth, FA(FM(0.5f, nx), 0.5f), FA(FM(0.5f, ny), 0.5f));
    sum += (nz < FA(sample_depth, 0.0000001f)) ? 1.0f : 0.0f;
  }
  ((unsigned char*)out.p)[(size_t)py * out.pitch + px] = (unsigned char)to_unorm(sum / 16.0f, 255.f);
}

struct SsrK {
  Mat4 normal_mat;
  CamK cam;
};

// The pass's depth sampler: NEAREST, clamp-to-border (opaque black = 0) in U, clamp-to-edge in V (ssr.cpp:22-28).
__device__ __forceinline__ float depth_nearest(const Mip& m, float u, float v) {
  const float fx = floorf(FM(u, (float)m.w)), fy = floorf(FM(v, (float)m.h));
  if (!(fx >= 0.0f && fx < (float)m.w)) return 0.0f;
  int y = (fy >= 0.0f && fy < 2147483648.f) ? (int)fy : 0;             // negative, NaN and out-of-int-range rows clamp to row 0 like the host's cast
  y = min(y, m.h - 1);
  return d24_to_float(ld32(m, (int)fx, y));
}

__device__ __forceinline__ float smoothstep_g(float e0, float e1, float x) {
  float t = gclamp(FS(x, e0) / FS(e1, e0), 0.0f, 1.0f);
  return FM(FM(t, t), FS(3.0f, FM(2.0f, t)));
}

__device__ __forceinline__ float4 unpack_rgba16f(uint2 k) {
  float2 a = unpack_rg16f(k.x), b = unpack_rg16f(k.y);
  return make_float4(a.x, a.y, b.x, b.y);
}
__device__ __forceinline__ float4 sample_rgba16f(const Mip& m, float u, float v) {
  Tap t = make_tap(u, v, m.w, m.h);
  float4 a = unpack_rgba16f(ld64(m, t.x0, t.y0)), b = unpack_rgba16f(ld64(m, t.x1, t.y0));
  float4 c = unpack_rgba16f(ld64(m, t.x0, t.y1)), d = unpack_rgba16f(ld64(m, t.x1, t.y1));
  return make_float4(bilerp(a.x, b.x, c.x, d.x, t.wx, t.wy), bilerp(a.y, b.y, c.y, d.y, t.wx, t.wy),
                     bilerp(a.z, b.z, c.z, d.z, t.wx, t.wy), bilerp(a.w, b.w, c.w, d.w, t.wx, t.wy));
}

// FRAME16F: the lit frame is RGBA16F (the TAA target) instead of RGBA8_SRGB (the shading output).
template <bool LAYERED, bool FRAME16F>
__global__ void __launch_bounds__(128) ssr_legacy_kernel(SsrK k0, const SsrK* __restrict__ kv, LaunchX lx, Dim dim, PyramidView P, Mip normal, Mip frame,
                                                         Mip material, Mip out, const float* __restrict__ srgb) {
  SsrK kl;
  unsigned long long poff = 0;
  if (LAYERED) {
    const int z = blockIdx.z;
    kl = kv[z];
    poff = (unsigned long long)z * P.mip[0].ls;
    normal = layer_of(normal, z); frame = layer_of(frame, z); material = layer_of(material, z); out = layer_of(out, z);
  }
  const SsrK& k = LAYERED ? kl : k0;
  const int px = blockIdx.x * 16 + (threadIdx.x & 15), py = VKR_BLOCK_Y * 8 + (threadIdx.x >> 4);
  if (px >= out.w || py >= out.h) return;
  const float su = pixel_uv(px, dim.w, dim.rw), sv = pixel_uv(py, dim.h, dim.rh);
  const CamK cam = k.cam;
  Mip d0 = P.mip[0];
  d0.p += poff;
  float4 result = make_float4(0.f, 0.f, 0.f, 0.f);
  do {
    const float tsx = (float)frame.w, tsy = (float)frame.h;                                                  // shader.frag:29-30
    const float au = FA(floorf(FM(su, tsx)) / tsx, 0.5f / tsx), av = FA(floorf(FM(sv, tsy)) / tsy, 0.5f / tsy);
    const float roughness = sample_srgb8(material, su, sv, srgb).y;
    const float pixel_depth = depth_nearest(d0, au, av);
    const f3 pixel_normal = norm3(mul_m4(k.normal_mat, sample_gbuffer_normal(normal, au, av), 0.f));
    const f3 view_vec = reconstruct_view_vec(au, av, pixel_depth, cam);
    const f3 R = reflect3(view_vec, pixel_normal);
    const f3 start = project_view_vec(view_vec + 0.0005f * pixel_normal, cam);
    const f3 pp = project_view_vec(view_vec + R, cam);
    const f3 delta = norm3(pp - start);
    if (fabsf(delta.z) < 0.0000001f) break;                                                                   // :49-52
    float t_bound = FS(1.f, start.z) / delta.z;
    const float u_bound = gmax(FS(1.f, start.x) / delta.x, -start.x / delta.x);
    const float v_bound = gmax(FS(1.f, start.y) / delta.y, -start.y / delta.y);
    t_bound = gmin(t_bound, gmin(u_bound, v_bound));
    const f3 end = start + t_bound * delta;

    // hierarchical_raymarch(depth_tex, start, end - start, 0, 100), screen_trace.glsl:51-100
    const f3 origin = start, direction = end - start;
    const f3 inv_direction = mk3(direction.x != 0 ? 1.0f / direction.x : MAX_T_FLOAT, direction.y != 0 ? 1.0f / direction.y : MAX_T_FLOAT,
                                 direction.z != 0 ? 1.0f / direction.z : MAX_T_FLOAT);
    int current_mip = 0;
    const float ssx = (float)d0.w, ssy = (float)d0.h;
    float resx = ssx, resy = ssy;
    float rinvx = 1.0f / resx, rinvy = 1.0f / resy;
    float uox = 0.005f / ssx, uoy = 0.005f / ssy;
    uox = direction.x < 0 ? -uox : uox;
    uoy = direction.y < 0 ? -uoy : uoy;
    const float fox = direction.x < 0 ? 0.f : 1.f, foy = direction.y < 0 ? 0.f : 1.f;
    float current_t;
    f3 position;
    {
      float xp = FA(FM(FA(floorf(FM(resx, origin.x)), fox), rinvx), uox);
      float yp = FA(FM(FA(floorf(FM(resy, origin.y)), foy), rinvy), uoy);
      float tx = FM(FS(xp, origin.x), inv_direction.x), ty = FM(FS(yp, origin.y), inv_direction.y);
      current_t = gmin(tx, ty);
      position = origin + current_t * direction;
    }
    int i = 0;
    while (i < 100 && current_mip >= 0) {
      float mpx = FM(resx, position.x), mpy = FM(resy, position.y);
      float surface_z = fetch_pyramid(P, poff, (int)mpx, (int)mpy, current_mip);
      bool skipped = advance_ray(origin, direction, inv_direction, mpx, mpy, rinvx, rinvy, fox, foy, uox, uoy, surface_z, position, current_t);
      current_mip += skipped ? 1 : -1;
      float f = skipped ? 0.5f : 2.0f, g = skipped ? 2.0f : 0.5f;
      resx *= f; resy *= f; rinvx *= g; rinvy *= g;
      ++i;
    }
    const f3 out_ray = position;                                                                              // valid_hit = (i <= 100): always

    if (fabsf(FS(out_ray.x, start.x)) < 2.0f / tsx && fabsf(FS(out_ray.y, start.y)) < 2.0f / tsy) break;      // :70-76
    const f3 hit_normal = mul_m4(k.normal_mat, sample_gbuffer_normal(normal, out_ray.x, out_ray.y), 0.f);
    if (dot3(hit_normal, R) > 0) break;                                                                       // :81-84
    const float hit_depth = depth_nearest(d0, out_ray.x, out_ray.y);
    if (out_ray.z > FA(hit_depth, 0.0001f)) break;                                                            // :86-90
    const float fovx = FM(0.05f, tsy / tsx), fovy = 0.05f;
    const float bx = FM(smoothstep_g(0.f, fovx, out_ray.x), FS(1.f, smoothstep_g(FS(1.f, fovx), 1.f, out_ray.x)));
    const float by = FM(smoothstep_g(0.f, fovy, out_ray.y), FS(1.f, smoothstep_g(FS(1.f, fovy), 1.f, out_ray.y)));
    const float coef = FM(bx, by);
    const float4 c = FRAME16F ? sample_rgba16f(frame, out_ray.x, out_ray.y) : sample_srgb8(frame, out_ray.x, out_ray.y, srgb);
    const float D = DistributionGGX(pixel_normal, pixel_normal, roughness);
    const float nr = gmax(dot3(pixel_normal, R), 0.0f);
    result = make_float4(FM(FM(FM(coef, c.x), D), nr), FM(FM(FM(coef, c.y), D), nr), FM(FM(FM(coef, c.z), D), nr), FM(FM(FM(coef, c.w), D), nr));
  } while (false);
  *((uint32_t*)(out.p + (size_t)py * out.pitch) + px) = pack_rgba8unorm(result.x, result.y, result.z, result.w);
}

}  // namespace vkr
using namespace vkr;

static SsaoK make_ssao_k(const VkrSsaoParams* p) {
  SsaoK k;
  k.projection = make_mat(p->projection);
  k.cam = make_cam(p->fovy, p->aspect, p->znear, p->zfar);
  return k;
}

extern "C" int vkr_ssao_ex(const VkrSsaoParams* p, const VkrImage* depth, const VkrImage* out_occlusion, const VkrLaunchExt* ext, void* stream) {
  if (!p || !image_ok(depth, VKR_FORMAT_D24_UNORM_S8_UINT) || !image_ok(out_occlusion, VKR_FORMAT_R8_UNORM)) {
    set_error("ssao: null argument or unexpected format");
    return VKR_ERR_INVALID_ARGUMENT;
  }
  ExtInfo x;
  int rc = resolve_ext(ext, out_occlusion->height, 8, &x, "ssao");
  if (rc) return rc;
  cudaStream_t st = (cudaStream_t)stream;
  SsaoK k = make_ssao_k(p);
  const SsaoK* kv = nullptr;
  if (x.layers > 1) {
    if (!x.prepared()) {
      std::vector<SsaoK> ks(x.layers);
      for (int i = 0; i < x.layers; i++) ks[i] = make_ssao_k(p + i);
      if ((rc = upload_params(ks.data(), sizeof(SsaoK), x.layers, x.scratch(), st, "ssao"))) return rc;
    }
    kv = (const SsaoK*)x.scratch() + x.first();
    if (x.prepare_only()) return VKR_OK;
  }
  SsaoSamples S;
  for (int i = 0; i < 16; i++) S.s[i] = make_float4(p->samples[i][0], p->samples[i][1], p->samples[i][2], 0.f);
  dim3 grid((out_occlusion->width + 31) / 32, x.yblocks, x.layers);
  const Dim dim = make_dim(out_occlusion->width, out_occlusion->height);
  auto kern = kv ? ssao_kernel<true> : ssao_kernel<false>;
  kern<<<grid, 256, 0, st>>>(k, kv, S, LaunchX{x.yb0}, dim, make_mip(depth, 0, x, 0), make_mip(out_occlusion, 0, x, 1));
  return check_launch("ssao_kernel");
}

extern "C" int vkr_ssao(const VkrSsaoParams* p, const VkrImage* depth, const VkrImage* out_occlusion, void* stream) {
  return vkr_ssao_ex(p, depth, out_occlusion, nullptr, stream);
}

static SsrK make_ssr_k(const VkrSsrParams* p) {
  SsrK k;
  k.normal_mat = make_mat(p->normal_mat);
  k.cam = make_cam(p->fovy, p->aspect, p->znear, p->zfar);
  return k;
}

extern "C" int vkr_ssr_legacy_ex(const VkrSsrParams* p, const VkrImage* depth, const VkrImage* normal, const VkrImage* frame,
                                 const VkrImage* material, const VkrImage* out_reflection, const VkrLaunchExt* ext, void* stream) {
  if (!p || !image_ok(depth, VKR_FORMAT_D24_UNORM_S8_UINT) || !image_ok(normal, VKR_FORMAT_R16G16_UNORM) ||
      !(image_ok(frame, VKR_FORMAT_R8G8B8A8_SRGB) || image_ok(frame, VKR_FORMAT_R16G16B16A16_SFLOAT)) ||
      !image_ok(material, VKR_FORMAT_R8G8B8A8_SRGB) || !image_ok(out_reflection, VKR_FORMAT_R8G8B8A8_UNORM)) {
    set_error("ssr_legacy: null argument or unexpected format");
    return VKR_ERR_INVALID_ARGUMENT;
  }
  if (frame->base == out_reflection->base) { set_error("ssr_legacy: frame and output alias"); return VKR_ERR_INVALID_ARGUMENT; }
  ExtInfo x;
  int rc = resolve_ext(ext, out_reflection->height, 8, &x, "ssr_legacy");
  if (rc) return rc;
  cudaStream_t st = (cudaStream_t)stream;
  SsrK k = make_ssr_k(p);
  const SsrK* kv = nullptr;
  if (x.layers > 1) {
    if (!x.prepared()) {
      std::vector<SsrK> ks(x.layers);
      for (int i = 0; i < x.layers; i++) ks[i] = make_ssr_k(p + i);
      if ((rc = upload_params(ks.data(), sizeof(SsrK), x.layers, x.scratch(), st, "ssr_legacy"))) return rc;
    }
    kv = (const SsrK*)x.scratch() + x.first();
    if (x.prepare_only()) return VKR_OK;
  }
  PyramidView P;                                   // every mip of the depth image (ssr.cpp:45 samples the whole chain)
  P.count = (int)depth->mip_count;
  for (int i = 0; i < VKR_MAX_MIPS; i++) P.mip[i] = (i < P.count) ? make_mip(depth, i, x, 0) : Mip{nullptr, 0, 0, 0, 0};
  const float* srgb = srgb_decode_table_device();
  if (!srgb) { set_error("ssr_legacy: sRGB table upload failed"); return VKR_ERR_CUDA; }
  dim3 grid((out_reflection->width + 15) / 16, x.yblocks, x.layers);
  const Dim dim = make_dim(out_reflection->width, out_reflection->height);
  const bool f16 = frame->format == VKR_FORMAT_R16G16B16A16_SFLOAT;
  auto kern = f16 ? (kv ? ssr_legacy_kernel<true, true> : ssr_legacy_kernel<false, true>) : (kv ? ssr_legacy_kernel<true, false> : ssr_legacy_kernel<false, false>);
  kern<<<grid, 128, 0, st>>>(k, kv, LaunchX{x.yb0}, dim, P, make_mip(normal, 0, x, 1), make_mip(frame, 0, x, 2), make_mip(material, 0, x, 3),
                             make_mip(out_reflection, 0, x, 4), srgb);
  return check_launch("ssr_legacy_kernel");
}

extern "C" int vkr_ssr_legacy(const VkrSsrParams* p, const VkrImage* depth, const VkrImage* normal, const VkrImage* frame,
                              const VkrImage* material, const VkrImage* out_reflection, void* stream) {
  return vkr_ssr_legacy_ex(p, depth, normal, frame, material, out_reflection, nullptr, stream);
}
