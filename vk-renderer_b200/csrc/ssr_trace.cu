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
namespace vkr {

struct TraceK {
  Mat4 normal_mat;
  CamK cam;
  uint32_t frame_random;
  float max_roughness;
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
                                                         const float4* __restrict__ halton, Mip pdf, Mip out_rays, Mip out_occl, const float* __restrict__ srgb) {
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
using namespace vkr;

static TraceK make_trace_k(const VkrTraceParams* p) {
  TraceK k;
  k.normal_mat = make_mat(p->normal_mat);
  k.cam = make_cam(p->fovy, p->aspect, p->znear, p->zfar);
  k.frame_random = p->frame_random; k.max_roughness = p->max_roughness;
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
  int rc = resolve_ext(ext, out_rays->height, 8, &x, "sssr_trace");
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
  auto kern = kv ? sssr_trace_kernel<true> : sssr_trace_kernel<false>;
  kern<<<grid, 128, 0, st>>>(k, kv, LaunchX{x.yb0}, P, make_mip(dnormal, 0, x, 1), make_mip(material, 0, x, 2), (const float4*)halton_vec4x128,
                             make_mip(pdf_lut, 0, x, 3), make_mip(out_rays, 0, x, 4), make_mip(out_occlusion, 0, x, 5), srgb);
  return check_launch("sssr_trace_kernel");
}

extern "C" int vkr_sssr_trace(const VkrTraceParams* p, const VkrImage* depth, const VkrImage* dnormal, const VkrImage* material,
                              const float* halton_vec4x128, const VkrImage* pdf_lut, const VkrImage* out_rays,
                              const VkrImage* out_occlusion, void* stream) {
  return vkr_sssr_trace_ex(p, depth, dnormal, material, halton_vec4x128, pdf_lut, out_rays, out_occlusion, nullptr, stream);
}
