// screen_trace.cu — ScreenSpaceTrace: tangent-space AO + one-bounce colour gather, 4x4 depth-aware filter, static
// temporal accumulation.  Replaces tasks "ScreenTrace", "ScreenTraceFilter", "ScreenTraceAccumulate"
// (reference src/screen_trace.cpp:23-180, shaders/screen_trace/{trace,filter,accumulate}.comp).
//
// Defined here where the reference is undefined (same choices as oracle/passes_screen_trace.cpp):
//   * the tile's shared-memory exchange (trace.comp:305-306 has memoryBarrierShared() but no barrier(), and sky pixels
//     return before writing their slot) behaves as if every invocation wrote first — sky pixels "no hit" — then gathered;
//   * only floor(W/8) x floor(H/8) tiles (trace) / floor(W/8)*8 x floor(H/4)*4 pixels (filter, accumulate) are written, and
//     uv = pixel / covered extent, as the reference's dispatch sizes and gl_NumWorkGroups arithmetic give.
// Everything that feeds a break / hit test is unfused IEEE; the three direction angles go through double sin/cos rounded
// once (the oracle does the same), so both sides march from identical bits.
#include <vector>
#include "brdf.cuh"

namespace vkr {

struct StK {
  Mat4 normal_mat;
  CamK cam;
  float random_offset, angle_offset;
};

#define ST_TILE 8
#define ST_MAX_THIKNESS 0.2f
#define ST_SAMPLES 20
#define ST_FAR_SAMPLES 8

__device__ __forceinline__ float st_rand(float u, float v) {                                       // trace.comp:42-44
  float d = FA(FM(u, 12.9898f), FM(v, 78.233f));
  float s = (float)sin((double)d);
  float p = FM(s, 43758.5453f);
  return FS(p, floorf(p));
}
__device__ __forceinline__ float sin_r(float x) { return (float)sin((double)x); }
__device__ __forceinline__ float cos_r(float x) { return (float)cos((double)x); }

__device__ __forceinline__ f3 st_sample_normal(const Mat4& M, const Mip& normal, float u, float v) {   // :46-48
  return norm3(mul_m4(M, sample_gbuffer_normal(normal, u, v), 0.f));
}

__device__ __forceinline__ float4 unpack_rgba16f(uint2 k) {
  float2 a = unpack_rg16f(k.x), b = unpack_rg16f(k.y);
  return make_float4(a.x, a.y, b.x, b.y);
}
__device__ __forceinline__ f3 sample_rgb16f(const Mip& m, float u, float v) {
  Tap t = make_tap(u, v, m.w, m.h);
  float4 a = unpack_rgba16f(ld64(m, t.x0, t.y0)), b = unpack_rgba16f(ld64(m, t.x1, t.y0));
  float4 c = unpack_rgba16f(ld64(m, t.x0, t.y1)), d = unpack_rgba16f(ld64(m, t.x1, t.y1));
  return mk3(bilerp(a.x, b.x, c.x, d.x, t.wx, t.wy), bilerp(a.y, b.y, c.y, d.y, t.wx, t.wy), bilerp(a.z, b.z, c.z, d.z, t.wx, t.wy));
}

// One block = 4 tiles of 8x8 side by side (32 x 8 threads); hits are exchanged inside each tile only.
template <bool LAYERED, bool COLOR16F>
__global__ void __launch_bounds__(256) screen_trace_kernel(StK k0, const StK* __restrict__ kv, LaunchX lx, int cw, int ch, Mip depth, Mip normal, Mip color,
                                                           Mip material, Mip out, const float* __restrict__ srgb) {
  __shared__ float4 s_hit[ST_TILE][32];     // xyz = RAY_HIT_POS, w unused
  __shared__ float4 s_col[ST_TILE][32];     // xyz = RAY_HIT_COLOR
  StK kl;
  if (LAYERED) {
    const int z = blockIdx.z;
    kl = kv[z];
    depth = layer_of(depth, z); normal = layer_of(normal, z); color = layer_of(color, z); material = layer_of(material, z); out = layer_of(out, z);
  }
  const StK& k = LAYERED ? kl : k0;
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
  const int px = blockIdx.x * 32 + tx, py = VKR_BLOCK_Y * ST_TILE + ty;
  const bool covered = px < cw && py < ch;
  const CamK cam = k.cam;
  bool sky = true;
  float ao = 0.f, su = 0.f, sv = 0.f;
  f3 camera_pos = mk3(0, 0, 0), camera_normal = mk3(0, 0, 0);
  float4 hit = make_float4(-1.f, -1.f, -1.f, 0.f), hcol = make_float4(0.f, 0.f, 0.f, 0.f);
  if (covered) {
    su = (float)px / (float)cw; sv = (float)py / (float)ch;                                          // :31
    const float sz = sample_d24(depth, su, sv);
    if (!(sz >= 1.f)) {
      sky = false;
      camera_pos = reconstruct_view_vec(su, sv, sz, cam);
      camera_normal = st_sample_normal(k.normal_mat, normal, su, sv);
      camera_pos = camera_pos + 1e-6f * camera_normal;
      f3 tangent, bitangent;                                                                          // calc_tangent_space :220-232
      if (fabsf(camera_normal.z) > 0.0f) {
        float kk = sqrtf(FA(FM(camera_normal.y, camera_normal.y), FM(camera_normal.z, camera_normal.z)));
        tangent = mk3(0.0f, -camera_normal.z / kk, camera_normal.y / kk);
      } else {
        float kk = sqrtf(FA(FM(camera_normal.x, camera_normal.x), FM(camera_normal.y, camera_normal.y)));
        tangent = mk3(camera_normal.y / kk, -camera_normal.x / kk, 0.0f);
      }
      bitangent = cross3(camera_normal, tangent);
      const float base_angle = FA(FM(1.0f / 16.0f, (float)((((px + py) & 3) << 2) + (px & 3))), k.angle_offset);
      const float normal_angle = FM(VKR_PI / 2, st_rand(FA(su, k.random_offset), FA(sv, 0.f)));
      const float rad = gmin(200.0f / len3(camera_pos), 32.0f);
      const float radius = gmax(rad / (float)depth.w, rad / (float)depth.h);
      const float angle = FM(2 * VKR_PI, FA(base_angle, 0.0f));
      const f3 dir = norm3(cos_r(angle) * tangent + sin_r(angle) * bitangent + camera_normal * sin_r(normal_angle));
      f3 sd = project_view_vec(camera_pos + dir, cam) - mk3(su, sv, sz);
      const float l2 = sqrtf(FA(FM(sd.x, sd.x), FM(sd.y, sd.y)));
      sd = mk3(sd.x / l2, sd.y / l2, sd.z / l2) * radius;
      bool ray_hit = false;
      f3 hit_pos = mk3(0, 0, 0);
      float h_cos = 0.f, previous_z = camera_pos.z;
      for (int i = 0; i < ST_SAMPLES; i++) {
        const float f = (float)i / (float)ST_SAMPLES;
        const f3 tc = mk3(su, sv, sz) + f * sd;
        const float tc_depth = sample_d24(depth, tc.x, tc.y);
        const f3 cs = reconstruct_view_vec(tc.x, tc.y, tc_depth, cam);
        if (tc.x < 0 || tc.x > 1 || tc.y < 0 || tc.y > 1 || cs.z > FA(previous_z, ST_MAX_THIKNESS)) break;
        if (!ray_hit && FS(tc.z, 1e-6f) > tc_depth) { hit_pos = tc; ray_hit = true; }
        h_cos = gmax(h_cos, dot3(camera_normal, norm3(cs - camera_pos)));
        previous_z = cs.z;
      }
      h_cos = gmin(h_cos, 1.0f);
      const float h = acosf(h_cos);
      ao = FM(0.25f, FS(1.f, cosf(FM(2.f, h))));
      const f3 start_ray = mk3(su, sv, sz) + sd;                                                     // :289-304
      sd = sd * 2.0f;
      for (int i = 0; i < ST_FAR_SAMPLES; i++) {
        const float f = (float)i / (float)ST_FAR_SAMPLES;
        const f3 tc = start_ray + f * sd;
        const float tc_depth = sample_d24(depth, tc.x, tc.y);
        const float camera_z = linearize_depth2(tc_depth, cam.znear, cam.zfar);
        if (tc.x < 0 || tc.x > 1 || tc.y < 0 || tc.y > 1 || camera_z > FA(previous_z, 0.1f)) break;
        if (!ray_hit && FS(tc.z, 1e-6f) > tc_depth) { hit_pos = tc; ray_hit = true; }
        previous_z = camera_z;
      }
      if (ray_hit) {
        const f3 hit_normal = st_sample_normal(k.normal_mat, normal, hit_pos.x, hit_pos.y);
        ray_hit = dot3(camera_normal, hit_normal) < 0;
      }
      if (ray_hit) {
        hit = make_float4(hit_pos.x, hit_pos.y, hit_pos.z, 0.f);
        f3 c;
        if (COLOR16F) c = sample_rgb16f(color, hit_pos.x, hit_pos.y);
        else { float4 c4 = sample_srgb8(color, hit_pos.x, hit_pos.y, srgb); c = mk3(c4.x, c4.y, c4.z); }
        hcol = make_float4(c.x, c.y, c.z, 0.f);
      }
    }
  }
  s_hit[ty][tx] = hit;
  s_col[ty][tx] = hcol;
  __syncthreads();
  if (!covered) return;
  uint2* dst = (uint2*)(out.p + (size_t)py * out.pitch) + px;
  if (sky) { *dst = make_uint2(pack_half2(0.f, 0.f), pack_half2(0.f, 1.f)); return; }                // :240-242
  const f3 W0 = -norm3(camera_pos);                                                                   // :308-333
  float weight = 0.f;
  const float roughness = sample_srgb8(material, su, sv, srgb).y;
  f3 accum = mk3(0, 0, 0);
  const int x0 = tx & ~7, lxx = tx & 7;
  for (int x = lxx - 1; x <= lxx + 1; x++)
    for (int y = ty - 1; y <= ty + 1; y++)
      if (x >= 0 && x < ST_TILE && y >= 0 && y < ST_TILE) {
        const float4 rh = s_hit[y][x0 + x];
        if (rh.z >= 0) {
          const f3 chp = reconstruct_view_vec(rh.x, rh.y, rh.z, cam);
          const f3 L = norm3(chp - camera_pos);
          const f3 H = norm3(W0 + L);
          const float w = FM(DistributionGGX(camera_normal, H, roughness), gmax(dot3(camera_normal, L), 0.0f));
          weight = FA(weight, w);
          const float4 hc = s_col[y][x0 + x];
          accum = accum + mk3(hc.x, hc.y, hc.z) * w;
        }
      }
  const f3 rgb = (weight > 0) ? mk3(accum.x / weight, accum.y / weight, accum.z / weight) : mk3(0, 0, 0);
  *dst = make_uint2(pack_half2(rgb.x, rgb.y), pack_half2(rgb.z, FM(ao, 2.0f)));
}

// filter.comp:13-39 (the sky store at :19-21 does not return: the weighted mean overwrites it)
template <bool LAYERED>
__global__ void __launch_bounds__(256) screen_trace_filter_kernel(float znear, float zfar, LaunchX lx, int cw, int ch, Mip raw, Mip depth, Mip out) {
  if (LAYERED) { const int z = blockIdx.z; raw = layer_of(raw, z); depth = layer_of(depth, z); out = layer_of(out, z); }
  const int px = blockIdx.x * 32 + (threadIdx.x & 31), py = VKR_BLOCK_Y * 8 + (threadIdx.x >> 5);
  if (px >= cw || py >= ch) return;
  const float linear_depth = linearize_depth2(fetch_d24(depth, px, py), znear, zfar);
  const float scale = FM(linear_depth, 0.1f);
  float weight_sum = 0.f;
  float4 sum = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
  for (int x = 0; x < 4; x++)
#pragma unroll
    for (int y = 0; y < 4; y++) {
      const int sx = px + x - 2, sy = py + y - 2;
      const float sampled = linearize_depth2(fetch_d24(depth, sx, sy), znear, zfar);
      const float weight = gmax(0.0f, FS(1.0f, fabsf(FS(sampled, linear_depth)) / scale));
      weight_sum = FA(weight_sum, weight);
      const float4 r = in_range(raw, sx, sy) ? unpack_rgba16f(ld64(raw, sx, sy)) : make_float4(0.f, 0.f, 0.f, 0.f);
      sum = make_float4(FA(sum.x, FM(weight, r.x)), FA(sum.y, FM(weight, r.y)), FA(sum.z, FM(weight, r.z)), FA(sum.w, FM(weight, r.w)));
    }
  *((uint2*)(out.p + (size_t)py * out.pitch) + px) =
      make_uint2(pack_half2(sum.x / weight_sum, sum.y / weight_sum), pack_half2(sum.z / weight_sum, sum.w / weight_sum));
}

// accumulate.comp:20-40
template <bool LAYERED>
__global__ void __launch_bounds__(256) screen_trace_accumulate_kernel(CamK cam, LaunchX lx, int cw, int ch, Mip depth, Mip prev_depth, Mip filtered, Mip accum) {
  if (LAYERED) { const int z = blockIdx.z; depth = layer_of(depth, z); prev_depth = layer_of(prev_depth, z); filtered = layer_of(filtered, z); accum = layer_of(accum, z); }
  const int px = blockIdx.x * 32 + (threadIdx.x & 31), py = VKR_BLOCK_Y * 8 + (threadIdx.x >> 5);
  if (px >= cw || py >= ch) return;
  const float su = (float)px / (float)cw, sv = (float)py / (float)ch;
  const float4 new_sum = in_range(filtered, px, py) ? unpack_rgba16f(ld64(filtered, px, py)) : make_float4(0.f, 0.f, 0.f, 0.f);
  const f3 cur_view = reconstruct_view_vec(su, sv, fetch_d24(depth, px, py), cam);
  float4 out_sum = new_sum;
  const float sampled_depth = fetch_d24(prev_depth, px, py);
  uint2* dst = (uint2*)(accum.p + (size_t)py * accum.pitch) + px;
  const float4 sampled_sum = unpack_rgba16f(*dst);
  const float sampled_z = linearize_depth2(sampled_depth, cam.znear, cam.zfar);
  const float delta = fabsf(FS(sampled_z, cur_view.z));
  if (delta < 1e-6f && sampled_depth < 1.0f)
    out_sum = make_float4(gmix(sampled_sum.x, new_sum.x, 0.05f), gmix(sampled_sum.y, new_sum.y, 0.05f), gmix(sampled_sum.z, new_sum.z, 0.05f),
                          gmix(sampled_sum.w, new_sum.w, 0.05f));
  *dst = make_uint2(pack_half2(out_sum.x, out_sum.y), pack_half2(out_sum.z, out_sum.w));
}

}  // namespace vkr
using namespace vkr;

static StK make_st_k(const VkrScreenTraceParams* p) {
  StK k;
  k.normal_mat = make_mat(p->normal_mat);
  k.cam = make_cam(p->fovy, p->aspect, p->znear, p->zfar);
  k.random_offset = p->random_offset; k.angle_offset = p->angle_offset;
  return k;
}

extern "C" int vkr_screen_trace_main_ex(const VkrScreenTraceParams* p, const VkrImage* depth, const VkrImage* normal, const VkrImage* color,
                                        const VkrImage* material, const VkrImage* out_raw, const VkrLaunchExt* ext, void* stream) {
  if (!p || !image_ok(depth, VKR_FORMAT_D24_UNORM_S8_UINT) || !image_ok(normal, VKR_FORMAT_R16G16_UNORM) ||
      !(image_ok(color, VKR_FORMAT_R8G8B8A8_SRGB) || image_ok(color, VKR_FORMAT_R16G16B16A16_SFLOAT)) ||
      !image_ok(material, VKR_FORMAT_R8G8B8A8_SRGB) || !image_ok(out_raw, VKR_FORMAT_R16G16B16A16_SFLOAT)) {
    set_error("screen_trace_main: null argument or unexpected format");
    return VKR_ERR_INVALID_ARGUMENT;
  }
  if (color->base == out_raw->base) { set_error("screen_trace_main: color and output alias"); return VKR_ERR_INVALID_ARGUMENT; }
  const int cw = (int)(out_raw->width / 8 * 8), ch = (int)(out_raw->height / 8 * 8);      // cmd.dispatch(extent.width/8, extent.height/8, 1)
  if (!cw || !ch) return VKR_OK;
  ExtInfo x;
  int rc = resolve_ext(ext, (uint32_t)ch, ST_TILE, &x, "screen_trace_main");
  if (rc) return rc;
  cudaStream_t st = (cudaStream_t)stream;
  StK k = make_st_k(p);
  const StK* kv = nullptr;
  if (x.layers > 1) {
    if (!x.prepared()) {
      std::vector<StK> ks(x.layers);
      for (int i = 0; i < x.layers; i++) ks[i] = make_st_k(p + i);
      if ((rc = upload_params(ks.data(), sizeof(StK), x.layers, x.scratch(), st, "screen_trace_main"))) return rc;
    }
    kv = (const StK*)x.scratch() + x.first();
    if (x.prepare_only()) return VKR_OK;
  }
  const float* srgb = srgb_decode_table_device();
  if (!srgb) { set_error("screen_trace_main: sRGB table upload failed"); return VKR_ERR_CUDA; }
  dim3 grid((cw + 31) / 32, x.yblocks, x.layers);
  const bool f16 = color->format == VKR_FORMAT_R16G16B16A16_SFLOAT;
  auto kern = f16 ? (kv ? screen_trace_kernel<true, true> : screen_trace_kernel<false, true>) : (kv ? screen_trace_kernel<true, false> : screen_trace_kernel<false, false>);
  kern<<<grid, 256, 0, st>>>(k, kv, LaunchX{x.yb0}, cw, ch, make_mip(depth, 0, x, 0), make_mip(normal, 0, x, 1), make_mip(color, 0, x, 2),
                             make_mip(material, 0, x, 3), make_mip(out_raw, 0, x, 4), srgb);
  return check_launch("screen_trace_kernel");
}

extern "C" int vkr_screen_trace_main(const VkrScreenTraceParams* p, const VkrImage* depth, const VkrImage* normal, const VkrImage* color,
                                     const VkrImage* material, const VkrImage* out_raw, void* stream) {
  return vkr_screen_trace_main_ex(p, depth, normal, color, material, out_raw, nullptr, stream);
}

extern "C" int vkr_screen_trace_filter_ex(float znear, float zfar, const VkrImage* raw, const VkrImage* depth, const VkrImage* out_filtered,
                                          const VkrLaunchExt* ext, void* stream) {
  if (!image_ok(raw, VKR_FORMAT_R16G16B16A16_SFLOAT) || !image_ok(depth, VKR_FORMAT_D24_UNORM_S8_UINT) ||
      !image_ok(out_filtered, VKR_FORMAT_R16G16B16A16_SFLOAT)) {
    set_error("screen_trace_filter: null argument or unexpected format");
    return VKR_ERR_INVALID_ARGUMENT;
  }
  if (raw->base == out_filtered->base) { set_error("screen_trace_filter: raw and filtered alias"); return VKR_ERR_INVALID_ARGUMENT; }
  const int cw = (int)(out_filtered->width / 8 * 8), ch = (int)(out_filtered->height / 4 * 4);   // cmd.dispatch(extent.width/8, extent.height/4, 1)
  if (!cw || !ch) return VKR_OK;
  ExtInfo x;
  int rc = resolve_ext(ext, (uint32_t)ch, 8, &x, "screen_trace_filter");
  if (rc) return rc;
  if (x.prepare_only()) return VKR_OK;
  dim3 grid((cw + 31) / 32, x.yblocks, x.layers);
  auto kern = x.layers > 1 ? screen_trace_filter_kernel<true> : screen_trace_filter_kernel<false>;
  kern<<<grid, 256, 0, (cudaStream_t)stream>>>(znear, zfar, LaunchX{x.yb0}, cw, ch, make_mip(raw, 0, x, 0), make_mip(depth, 0, x, 1), make_mip(out_filtered, 0, x, 2));
  return check_launch("screen_trace_filter_kernel");
}

extern "C" int vkr_screen_trace_filter(float znear, float zfar, const VkrImage* raw, const VkrImage* depth, const VkrImage* out_filtered, void* stream) {
  return vkr_screen_trace_filter_ex(znear, zfar, raw, depth, out_filtered, nullptr, stream);
}

extern "C" int vkr_screen_trace_accumulate_ex(const float* fovy_aspect_znear_zfar, const VkrImage* depth, const VkrImage* prev_depth,
                                              const VkrImage* filtered, const VkrImage* accumulated, const VkrLaunchExt* ext, void* stream) {
  if (!fovy_aspect_znear_zfar || !image_ok(depth, VKR_FORMAT_D24_UNORM_S8_UINT) || !image_ok(prev_depth, VKR_FORMAT_D24_UNORM_S8_UINT) ||
      !image_ok(filtered, VKR_FORMAT_R16G16B16A16_SFLOAT) || !image_ok(accumulated, VKR_FORMAT_R16G16B16A16_SFLOAT)) {
    set_error("screen_trace_accumulate: null argument or unexpected format");
    return VKR_ERR_INVALID_ARGUMENT;
  }
  if (filtered->base == accumulated->base) { set_error("screen_trace_accumulate: filtered and accumulated alias"); return VKR_ERR_INVALID_ARGUMENT; }
  const int cw = (int)(accumulated->width / 8 * 8), ch = (int)(accumulated->height / 4 * 4);
  if (!cw || !ch) return VKR_OK;
  ExtInfo x;
  int rc = resolve_ext(ext, (uint32_t)ch, 8, &x, "screen_trace_accumulate");
  if (rc) return rc;
  if (x.prepare_only()) return VKR_OK;
  const CamK cam = make_cam(fovy_aspect_znear_zfar[0], fovy_aspect_znear_zfar[1], fovy_aspect_znear_zfar[2], fovy_aspect_znear_zfar[3]);
  dim3 grid((cw + 31) / 32, x.yblocks, x.layers);
  auto kern = x.layers > 1 ? screen_trace_accumulate_kernel<true> : screen_trace_accumulate_kernel<false>;
  kern<<<grid, 256, 0, (cudaStream_t)stream>>>(cam, LaunchX{x.yb0}, cw, ch, make_mip(depth, 0, x, 0), make_mip(prev_depth, 0, x, 1), make_mip(filtered, 0, x, 2),
                                               make_mip(accumulated, 0, x, 3));
  return check_launch("screen_trace_accumulate_kernel");
}

extern "C" int vkr_screen_trace_accumulate(const float* fovy_aspect_znear_zfar, const VkrImage* depth, const VkrImage* prev_depth,
                                           const VkrImage* filtered, const VkrImage* accumulated, void* stream) {
  return vkr_screen_trace_accumulate_ex(fovy_aspect_znear_zfar, depth, prev_depth, filtered, accumulated, nullptr, stream);
}
