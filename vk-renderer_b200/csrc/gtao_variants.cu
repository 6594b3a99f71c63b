// gtao_variants.cu — the GTAO variants beside the live compute path:
//   task "GTAO"               fragment-shader GTAO, 1 slice x 20 taps       (reference src/gtao.cpp:349-413, gtao/main.frag)
//   task "GTAO_deinterleave"  depth -> 4^n-layer R32F array                  (src/gtao.cpp:445-472, gtao_opt/deinterleave.comp)
//   task "GTAO_deinterleaved" the same slice integral on one array layer     (src/gtao.cpp:474-526, gtao_opt/main_deinterleaved.comp)
//   task "GTAO_reproject"     static-camera 5 % blend with the last frame    (src/gtao.cpp:241-284, gtao/reproject.comp)
// The dispatch-extent quirks of the reference are restated, not repaired (see oracle/passes_gtao_variants.cpp header).
// The march decides per tap whether to stop (thickness break), so everything up to the horizon cosine is unfused IEEE with the
// slice angle through double sin/cos rounded once, like the oracle; the closing integral is ordinary fp32.
#include <vector>
#include "common.cuh"

namespace vkr {

struct GvK {
  Mat4 normal_mat;
  CamK cam;
  float angle_offset;
};

#define GV_SAMPLES 20
#define GV_MAX_THIKNESS 0.1f

__device__ __forceinline__ float gv_sin(float x) { return (float)sin((double)x); }
__device__ __forceinline__ float gv_cos(float x) { return (float)cos((double)x); }

// R32F = the deinterleaved array layer, otherwise a D24 mip level.
template <bool R32F>
__device__ __forceinline__ float gv_depth(const Mip& m, float u, float v) { return R32F ? sample_r32f(m, u, v) : sample_d24(m, u, v); }

// main.frag:167-196 == main_deinterleaved.comp:87-116 with dirs_count = 1
template <bool R32F>
__device__ __forceinline__ float gtao_camera_space_v(const GvK& k, const Mip& depth, const Mip& normal, int px, int py, float su, float sv) {
  const float frag_depth = gv_depth<R32F>(depth, su, sv);
  if (frag_depth >= 1.f) return 1.f;
  const CamK cam = k.cam;
  const f3 camera_pos = reconstruct_view_vec(su, sv, frag_depth, cam);
  const f3 w0 = -norm3(camera_pos);
  const f3 camera_normal = norm3(mul_m4(k.normal_mat, sample_gbuffer_normal(normal, su, sv), 0.f));
  const float rad = gmin(200.0f / len3(camera_pos), 32.0f);
  const float rx = rad / (float)depth.w, ry = rad / (float)depth.h;
  const float base_angle = FA(FM(1.0f / 16.0f, (float)((((px + py) & 3) << 2) + (px & 3))), k.angle_offset);
  const float angle = FM(2 * VKR_PI, FA(base_angle, 0.0f));
  const float dx = FM(rx, gv_cos(angle)), dy = FM(ry, gv_sin(angle));
  const f3 sample_end_pos = reconstruct_view_vec(FA(su, dx), FA(sv, dy), frag_depth, cam);
  const f3 slice_normal = norm3(cross3(w0, -sample_end_pos));
  const f3 normal_projected = camera_normal - dot3(camera_normal, slice_normal) * slice_normal;
  const float n = FS(VKR_PI / 2.0f, acosf(dot3(norm3(normal_projected), norm3(sample_end_pos - camera_pos))));
  float h_cos = -1.0f, previous_z = camera_pos.z;                                                  // find_horizon, main.frag:69-93
  for (int i = 1; i <= GV_SAMPLES; i++) {
    const float f = (float)i / (float)GV_SAMPLES;
    const float tu = FA(su, FM(f, dx)), tv = FA(sv, FM(f, dy));
    const f3 sp = reconstruct_view_vec(tu, tv, gv_depth<R32F>(depth, tu, tv), cam);
    if (sp.z > FA(previous_z, GV_MAX_THIKNESS)) break;
    previous_z = sp.z;
    const float sample_cos = dot3(w0, norm3(sp - camera_pos));
    if (sample_cos > h_cos) h_cos = sample_cos;
  }
  float h = acosf(h_cos);
  h = gmin(FA(n, gmin(FS(h, n), VKR_PI / 2.0f)), h);
  const float sum = FM(FM(len3(normal_projected), 0.25f), gmax(FA(FA(-cosf(FS(FM(2.f, h), n)), cosf(n)), FM(FM(2.f, h), sinf(n))), 0.0f));
  return FM(2.f, sum);
}

__global__ void __launch_bounds__(256) gtao_graphics_kernel(GvK k, LaunchX lx, Dim dim, Mip depth, Mip normal, Mip raw) {
  const int px = blockIdx.x * 32 + (threadIdx.x & 31), py = VKR_BLOCK_Y * 8 + (threadIdx.x >> 5);
  if (px >= raw.w || py >= raw.h) return;
  const float su = pixel_uv(px, dim.w, dim.rw), sv = pixel_uv(py, dim.h, dim.rh);
  const float o = gtao_camera_space_v<false>(k, depth, normal, px, py, su, sv);
  *((uint2*)(raw.p + (size_t)py * raw.pitch) + px) = make_uint2(pack_half2(o, 0.f), pack_half2(0.f, 1.f));
}

// deinterleave.comp:10-21.  cw x ch = the dispatched extent (of the OUTPUT image, gtao.cpp:468).
__global__ void __launch_bounds__(256) gtao_deinterleave_kernel(int cw, int ch, Mip depth, Mip out, int pattern_step) {
  const int px = blockIdx.x * 32 + (threadIdx.x & 31), py = blockIdx.y * 8 + (threadIdx.x >> 5);
  if (px >= cw || py >= ch) return;
  const float d = fetch_d24(depth, px, py);
  const int ox = px >> pattern_step, oy = py >> pattern_step, mod = (1 << pattern_step) - 1;
  const int layer = ((py & mod) << pattern_step) + (px & mod);
  if (ox < out.w && oy < out.h) *((float*)(out.p + (size_t)layer * out.ls + (size_t)oy * out.pitch) + ox) = d;
}

// main_deinterleaved.comp:37-51: invocation (ix, iy) of the dispatch over raw's extent -> pixel 2^n * (ix, iy) + layer offset.
__global__ void __launch_bounds__(256) gtao_deinterleaved_kernel(GvK k, LaunchX lx, int cw, int ch, Mip depth_layer, Mip normal, Mip raw, int pattern_n, unsigned layer) {
  const int ix = blockIdx.x * 32 + (threadIdx.x & 31), iy = VKR_BLOCK_Y * 8 + (threadIdx.x >> 5);
  if (ix >= cw || iy >= ch) return;
  const int scale = 1 << pattern_n, mod = scale - 1;
  const int px = scale * ix + (int)(layer & mod), py = scale * iy + (int)((layer >> pattern_n) & mod);
  if (px >= raw.w || py >= raw.h) return;                                                          // the store would be dropped
  const float su = (float)px / (float)(scale * cw), sv = (float)py / (float)(scale * ch);
  const float o = gtao_camera_space_v<true>(k, depth_layer, normal, px, py, su, sv);
  *((uint2*)(raw.p + (size_t)py * raw.pitch) + px) = make_uint2(pack_half2(o, 0.f), 0u);
}

// reproject.comp:26-66, STATIC_REPROJECT
__global__ void __launch_bounds__(256) gtao_reproject_kernel(CamK cam, LaunchX lx, int cw, int ch, Mip depth, Mip prev_depth, Mip ao, Mip prev_ao, Mip out) {
  const int px = blockIdx.x * 32 + (threadIdx.x & 31), py = VKR_BLOCK_Y * 8 + (threadIdx.x >> 5);
  if (px >= cw || py >= ch) return;
  const float su = (float)px / (float)cw, sv = (float)py / (float)ch;
  const float new_ao = in_range(ao, px, py) ? __half2float(__ushort_as_half(ld16(ao, px, py))) : 0.f;
  const f3 cur_view = reconstruct_view_vec(su, sv, fetch_d24(depth, px, py), cam);
  float r = new_ao;
  const float sampled_depth = fetch_d24(prev_depth, px, py);
  const float sampled_ao = in_range(prev_ao, px, py) ? __half2float(__ushort_as_half(ld16(prev_ao, px, py))) : 0.f;
  const float sampled_z = linearize_depth2(sampled_depth, cam.znear, cam.zfar);
  if (fabsf(FS(sampled_z, cur_view.z)) < 1e-6f && sampled_depth < 1.0f) r = gmix(sampled_ao, new_ao, 0.05f);
  *((__half*)(out.p + (size_t)py * out.pitch) + px) = __float2half_rn(r);
}

}  // namespace vkr
using namespace vkr;

static GvK make_gv_k(const VkrGtaoParams* p) {
  GvK k;
  k.normal_mat = make_mat(p->normal_mat);
  k.cam = make_cam(p->fovy, p->aspect, p->znear, p->zfar);
  k.angle_offset = p->angle_offset;
  return k;
}

extern "C" int vkr_gtao_main_graphics(const VkrGtaoParams* p, const VkrImage* depth, const VkrImage* normal, const VkrImage* raw, uint32_t depth_lod,
                                      const VkrLaunchExt* ext, void* stream) {
  if (!p || !image_ok(depth, VKR_FORMAT_D24_UNORM_S8_UINT) || depth_lod >= depth->mip_count || !image_ok(normal, VKR_FORMAT_R16G16_UNORM) ||
      !image_ok(raw, VKR_FORMAT_R16G16B16A16_SFLOAT)) {
    set_error("gtao_main_graphics: null argument or unexpected format");
    return VKR_ERR_INVALID_ARGUMENT;
  }
  ExtInfo x;
  int rc = resolve_ext(ext, raw->height, 8, &x, "gtao_main_graphics");
  if (rc) return rc;
  if (x.layers > 1) { set_error("gtao_main_graphics: array layers are not supported"); return VKR_ERR_UNSUPPORTED; }
  dim3 grid((raw->width + 31) / 32, x.yblocks, 1);
  gtao_graphics_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(make_gv_k(p), LaunchX{x.yb0}, make_dim(raw->width, raw->height), make_mip(depth, depth_lod),
                                                                make_mip(normal, 0), make_mip(raw, 0));
  return check_launch("gtao_graphics_kernel");
}

extern "C" int vkr_gtao_deinterleave(const VkrImage* depth, uint32_t depth_lod, const VkrImage* out_array, uint64_t layer_stride, int pattern_n, void* stream) {
  if (!image_ok(depth, VKR_FORMAT_D24_UNORM_S8_UINT) || depth_lod >= depth->mip_count || !image_ok(out_array, VKR_FORMAT_R32_SFLOAT) || pattern_n < 0 ||
      pattern_n > 3 || !layer_stride) {
    set_error("gtao_deinterleave: null argument, unexpected format or pattern_n outside 0..3");
    return VKR_ERR_INVALID_ARGUMENT;
  }
  const int cw = (int)(out_array->width / 8 * 8), ch = (int)(out_array->height / 4 * 4);            // gtao.cpp:468
  if (!cw || !ch) return VKR_OK;
  Mip out = make_mip(out_array, 0);
  out.ls = layer_stride;
  dim3 grid((cw + 31) / 32, (ch + 7) / 8, 1);
  gtao_deinterleave_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(cw, ch, make_mip(depth, depth_lod), out, pattern_n);
  return check_launch("gtao_deinterleave_kernel");
}

extern "C" int vkr_gtao_main_deinterleaved(const VkrGtaoParams* p, const VkrImage* depth_array, uint64_t layer_stride, const VkrImage* normal,
                                           const VkrImage* raw, int pattern_n, uint32_t layer, const VkrLaunchExt* ext, void* stream) {
  if (!p || !image_ok(depth_array, VKR_FORMAT_R32_SFLOAT) || !image_ok(normal, VKR_FORMAT_R16G16_UNORM) || !image_ok(raw, VKR_FORMAT_R16G16B16A16_SFLOAT) ||
      pattern_n < 0 || pattern_n > 3 || layer >= (1u << (2 * pattern_n)) || (layer && !layer_stride)) {
    set_error("gtao_main_deinterleaved: null argument, unexpected format, or pattern_n / layer out of range");
    return VKR_ERR_INVALID_ARGUMENT;
  }
  const int cw = (int)(raw->width / 8 * 8), ch = (int)(raw->height / 4 * 4);                        // gtao.cpp:523
  if (!cw || !ch) return VKR_OK;
  ExtInfo x;
  int rc = resolve_ext(ext, (uint32_t)ch, 8, &x, "gtao_main_deinterleaved");
  if (rc) return rc;
  if (x.layers > 1) { set_error("gtao_main_deinterleaved: array layers are not supported"); return VKR_ERR_UNSUPPORTED; }
  Mip dl = make_mip(depth_array, 0);
  dl.p += (size_t)layer * layer_stride;
  dim3 grid((cw + 31) / 32, x.yblocks, 1);
  gtao_deinterleaved_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(make_gv_k(p), LaunchX{x.yb0}, cw, ch, dl, make_mip(normal, 0), make_mip(raw, 0), pattern_n, layer);
  return check_launch("gtao_deinterleaved_kernel");
}

extern "C" int vkr_gtao_reproject(const float* fovy_aspect_znear_zfar, const VkrImage* depth, const VkrImage* prev_depth, uint32_t depth_lod,
                                  const VkrImage* gtao, const VkrImage* prev_gtao, const VkrImage* out, const VkrLaunchExt* ext, void* stream) {
  if (!fovy_aspect_znear_zfar || !image_ok(depth, VKR_FORMAT_D24_UNORM_S8_UINT) || !image_ok(prev_depth, VKR_FORMAT_D24_UNORM_S8_UINT) ||
      depth_lod >= depth->mip_count || depth_lod >= prev_depth->mip_count || !image_ok(gtao, VKR_FORMAT_R16_SFLOAT) ||
      !image_ok(prev_gtao, VKR_FORMAT_R16_SFLOAT) || !image_ok(out, VKR_FORMAT_R16_SFLOAT)) {
    set_error("gtao_reproject: null argument or unexpected format");
    return VKR_ERR_INVALID_ARGUMENT;
  }
  if (out->base == gtao->base || out->base == prev_gtao->base) { set_error("gtao_reproject: output aliases an input"); return VKR_ERR_INVALID_ARGUMENT; }
  const int cw = (int)(out->width / 8 * 8), ch = (int)(out->height / 4 * 4);                        // gtao.cpp:282
  if (!cw || !ch) return VKR_OK;
  ExtInfo x;
  int rc = resolve_ext(ext, (uint32_t)ch, 8, &x, "gtao_reproject");
  if (rc) return rc;
  if (x.layers > 1) { set_error("gtao_reproject: array layers are not supported"); return VKR_ERR_UNSUPPORTED; }
  const CamK cam = make_cam(fovy_aspect_znear_zfar[0], fovy_aspect_znear_zfar[1], fovy_aspect_znear_zfar[2], fovy_aspect_znear_zfar[3]);
  dim3 grid((cw + 31) / 32, x.yblocks, 1);
  gtao_reproject_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(cam, LaunchX{x.yb0}, cw, ch, make_mip(depth, depth_lod), make_mip(prev_depth, depth_lod),
                                                                 make_mip(gtao, 0), make_mip(prev_gtao, 0), make_mip(out, 0));
  return check_launch("gtao_reproject_kernel");
}
