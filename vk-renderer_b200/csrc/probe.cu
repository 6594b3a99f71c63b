// probe.cu — the probe_renderer compute path:
//   task "Cubemap2Octahedral"  cube colour + radial distance -> octahedral colour / depth layer
//                              (reference src/probe_renderer.cpp:170-204, shaders/cube2oct/shader.comp, include/octahedral.glsl)
//   task "DownsampleProbe"     R16_UNORM min pyramid of a probe layer (src/probe_renderer.cpp:206-236, probe_downsample/shader.frag)
//   task "TraceProbe"          mirror ray through a grid of octahedral probes: per probe up to 4 octant segments, each a
//                              Hi-Z march of <= 25 cells in octahedral space (src/probe_renderer.cpp:316-385, trace_probe/shader.comp)
// Cube sampling follows the interpretation pinned in oracle/passes_probe.cpp (face table, bilinear clamped to the face).
// All three pick texels / branch on thresholds, so they are evaluated unfused like the oracle.
#include <vector>
#include "march.cuh"
#include "../../include/vkr_probe.h"

namespace vkr {

#define PR_ZNEAR 0.05f
#define PR_ZFAR 80.f

struct v2f { float x, y; };
__device__ __forceinline__ v2f mk2(float x, float y) { v2f r; r.x = x; r.y = y; return r; }
__device__ __forceinline__ float o_sign_nz(float k) { return (k >= 0.0f) ? 1.0f : -1.0f; }
__device__ __forceinline__ float sign_g(float x) { return x > 0.0f ? 1.0f : (x < 0.0f ? -1.0f : 0.0f); }

__device__ __forceinline__ f3 oct_unfold(float u, float v) {                                       // octahedral.glsl:38-45
  u = FM(2.f, FS(u, 0.5f)); v = FM(2.f, FS(v, 0.5f));
  f3 r = mk3(u, v, FS(FS(1.0f, fabsf(u)), fabsf(v)));
  if (r.z < 0.0f) {
    float x = FM(FS(1.0f, fabsf(r.y)), o_sign_nz(r.x)), y = FM(FS(1.0f, fabsf(r.x)), o_sign_nz(r.y));
    r.x = x; r.y = y;
  }
  return r;
}
__device__ __forceinline__ f3 oct_decode(float u, float v) { return norm3(oct_unfold(u, v)); }
__device__ __forceinline__ f3 oct_center(float u, float v) { f3 r = oct_unfold(u, v); return norm3(mk3(sign_g(r.x), sign_g(r.y), sign_g(r.z))); }
__device__ __forceinline__ v2f oct_encode(f3 v) {                                                   // :59-66
  float l1 = FA(FA(fabsf(v.x), fabsf(v.y)), fabsf(v.z));
  float inv = 1.0f / l1;
  float rx = FM(v.x, inv), ry = FM(v.y, inv);
  if (v.z < 0.0f) {
    float tx = FM(FS(1.0f, fabsf(ry)), o_sign_nz(rx)), ty = FM(FS(1.0f, fabsf(rx)), o_sign_nz(ry));
    rx = tx; ry = ty;
  }
  return mk2(FA(FM(0.5f, rx), 0.5f), FA(FM(0.5f, ry), 0.5f));
}
__device__ __forceinline__ float encode_oct_depth(float z, float n, float f) { return FA(f / FS(f, n), FM(f, n) / FM(-z, FS(f, n))); }
__device__ __forceinline__ float decode_oct_depth(float d, float n, float f) { return FM(-n, f) / FS(FM(d, FS(f, n)), f); }

__device__ __forceinline__ float half_at(const Mip& m, int x, int y) { return __half2float(__ushort_as_half(ld16(m, x, y))); }
__device__ __forceinline__ float sample_r16f(const Mip& m, float u, float v) {
  Tap t = make_tap(u, v, m.w, m.h);
  return bilerp(half_at(m, t.x0, t.y0), half_at(m, t.x1, t.y0), half_at(m, t.x0, t.y1), half_at(m, t.x1, t.y1), t.wx, t.wy);
}
__device__ __forceinline__ float sample_r16unorm(const Mip& m, float u, float v) {
  Tap t = make_tap(u, v, m.w, m.h);
  return bilerp(unorm16_to_float(ld16(m, t.x0, t.y0)), unorm16_to_float(ld16(m, t.x1, t.y0)), unorm16_to_float(ld16(m, t.x0, t.y1)),
                unorm16_to_float(ld16(m, t.x1, t.y1)), t.wx, t.wy);
}
__device__ __forceinline__ float4 sample_rgba8unorm(const Mip& m, float u, float v) {
  Tap t = make_tap(u, v, m.w, m.h);
  uint32_t a = ld32(m, t.x0, t.y0), b = ld32(m, t.x1, t.y0), c = ld32(m, t.x0, t.y1), d = ld32(m, t.x1, t.y1);
  float r[4];
#pragma unroll
  for (int i = 0; i < 4; i++)
    r[i] = bilerp(unorm8_to_float((a >> (8 * i)) & 255u), unorm8_to_float((b >> (8 * i)) & 255u), unorm8_to_float((c >> (8 * i)) & 255u),
                  unorm8_to_float((d >> (8 * i)) & 255u), t.wx, t.wy);
  return make_float4(r[0], r[1], r[2], r[3]);
}

// Face and (s, t) of a direction (Vulkan cube map face selection; ties z over y over x).
__device__ __forceinline__ int cube_face(f3 r, float& s, float& t) {
  const float ax = fabsf(r.x), ay = fabsf(r.y), az = fabsf(r.z);
  int face; float sc, tc, rc;
  if (az >= ax && az >= ay) { face = r.z < 0 ? 5 : 4; sc = r.z < 0 ? -r.x : r.x; tc = -r.y; rc = az; }
  else if (ay >= ax)        { face = r.y < 0 ? 3 : 2; sc = r.x; tc = r.y < 0 ? -r.z : r.z; rc = ay; }
  else                      { face = r.x < 0 ? 1 : 0; sc = r.x < 0 ? r.z : -r.z; tc = -r.y; rc = ax; }
  s = FA(FM(0.5f, sc / rc), 0.5f); t = FA(FM(0.5f, tc / rc), 0.5f);
  return face;
}

// cube2oct/shader.comp:14-33.  grid.z = probes converted in one launch (cube sets and target layers at fixed strides).
__global__ void __launch_bounds__(256) cube2oct_kernel(int cw, int ch, Mip cube_color, Mip cube_dist, unsigned long long cube_color_set, unsigned long long cube_dist_set,
                                                       Mip oct_color, Mip oct_depth, const float* __restrict__ srgb) {
  const int px = blockIdx.x * 32 + (threadIdx.x & 31), py = blockIdx.y * 8 + (threadIdx.x >> 5), z = blockIdx.z;
  if (px >= cw || py >= ch) return;
  cube_color.p += (size_t)z * cube_color_set; cube_dist.p += (size_t)z * cube_dist_set;
  oct_color = layer_of(oct_color, z); oct_depth = layer_of(oct_depth, z);
  const float su = (float)px / (float)cw, sv = (float)py / (float)ch;
  const f3 dir = oct_decode(su, sv);
  float s, t;
  const int face = cube_face(dir, s, t);
  const float4 col = sample_srgb8(layer_of(cube_color, face), s, t, srgb);
  const float dist = sample_r16f(layer_of(cube_dist, face), s, t);
  const f3 view_dir = dir * dist;
  const f3 front = oct_center(su, sv);
  const float depth = encode_oct_depth(gclamp(dot3(view_dir, front), PR_ZNEAR, PR_ZFAR), PR_ZNEAR, PR_ZFAR);
  *((uint32_t*)(oct_color.p + (size_t)py * oct_color.pitch) + px) = pack_rgba8unorm(col.x, col.y, col.z, col.w);
  *((uint16_t*)(oct_depth.p + (size_t)py * oct_depth.pitch) + px) = (uint16_t)to_unorm(depth, 65535.f);
}

// probe_downsample/shader.frag:8-23 — one mip from the previous one; grid.z = layers.
__global__ void __launch_bounds__(256) probe_downsample_kernel(Mip src, Mip dst, int w, int h) {
  const int px = blockIdx.x * 32 + (threadIdx.x & 31), py = blockIdx.y * 8 + (threadIdx.x >> 5);
  if (px >= w || py >= h) return;
  src = layer_of(src, blockIdx.z); dst = layer_of(dst, blockIdx.z);
  float min_depth = 1000.f;
#pragma unroll
  for (int i = 0; i < 4; i++) {
    const int qx = min(2 * px + (i >> 1), src.w), qy = min(2 * py + (i & 1), src.h);           // `min(pixel + off, size)`: index == size fetches 0
    const float d = in_range(src, qx, qy) ? unorm16_to_float(ld16(src, qx, qy)) : 0.0f;
    min_depth = gmin(d, min_depth);
  }
  *((uint16_t*)(dst.p + (size_t)py * dst.pitch) + px) = (uint16_t)to_unorm(min_depth, 65535.f);
}

}  // namespace vkr

namespace vkr {

struct ProbeK {
  Mat4 inverse_view;
  CamK cam;
  f3 probe_min, probe_max;
  uint32_t grid_size;
};

enum { TRACE_RESULT_MISS = 0, TRACE_RESULT_HIT = 1, TRACE_RESULT_UNKNOWN = 2 };
#define PR_TRACE_STEPS 25

struct ProbeArrays {
  PyramidView depth;                 // layer 0 of the R16_UNORM min-pyramid array; mip[0].ls = bytes between layers
  Mip color;                         // layer 0 of the RGBA8_UNORM array
};

__device__ __forceinline__ float fetch_probe(const PyramidView& P, unsigned long long layer_off, int x, int y, int lod) {
  if ((unsigned)lod >= (unsigned)P.count) return 0.0f;
  const Mip& m = P.mip[lod];
  if ((unsigned)x >= (unsigned)m.w || (unsigned)y >= (unsigned)m.h) return 0.0f;
  return unorm16_to_float(__ldg((const uint16_t*)(m.p + layer_off + (size_t)y * m.pitch) + x));
}

__device__ __forceinline__ void swap_min(float& a, float& b) { float t = gmin(a, b); b = gmax(a, b); a = t; }

// trace_probe/shader.comp:224-275 with the shader's own advance_ray (:191-218: t_min capped at 1)
__device__ __forceinline__ f3 probe_raymarch(const PyramidView& P, unsigned long long poff, f3 origin, f3 direction) {
  const f3 inv_direction = mk3(direction.x != 0 ? 1.0f / direction.x : MAX_T_FLOAT, direction.y != 0 ? 1.0f / direction.y : MAX_T_FLOAT,
                               direction.z != 0 ? 1.0f / direction.z : MAX_T_FLOAT);
  int current_mip = 0;
  const float ssx = (float)P.mip[0].w, ssy = (float)P.mip[0].h;
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
  while (i < PR_TRACE_STEPS && current_mip >= 0) {
    const float mpx = FM(resx, position.x), mpy = FM(resy, position.y);
    const float surface_z = fetch_probe(P, poff, (int)mpx, (int)mpy, current_mip);
    const float xp = FA(FM(FA(floorf(mpx), fox), rinvx), uox);
    const float yp = FA(FM(FA(floorf(mpy), foy), rinvy), uoy);
    const float tx = FM(FS(xp, origin.x), inv_direction.x), ty = FM(FS(yp, origin.y), inv_direction.y);
    float tz = FM(FS(surface_z, origin.z), inv_direction.z);
    tz = direction.z > 0 ? tz : MAX_T_FLOAT;
    const float t_min = gmin(gmin(gmin(tx, ty), tz), 1.0f);
    const bool above_surface = surface_z > position.z;
    const bool skipped = t_min != tz && above_surface;
    current_t = above_surface ? t_min : current_t;
    position = origin + current_t * direction;
    current_mip += skipped ? 1 : -1;
    const float f = skipped ? 0.5f : 2.0f, g = skipped ? 2.0f : 0.5f;
    resx *= f; resy *= f; rinvx *= g; rinvy *= g;
    ++i;
  }
  return position;
}

// :277-331
__device__ __forceinline__ int trace_segment_hi(const ProbeArrays& A, f3 ray_origin, f3 ray_dir, float t0, float t1, float& tmin, int probe_index, v2f& stop_tc) {
  const float RAY_EPS = 0.001f;
  f3 probe_start = ray_origin + ray_dir * FA(t0, RAY_EPS);
  const f3 probe_end = ray_origin + ray_dir * FS(t1, RAY_EPS);
  { const f3 d = probe_start - probe_end; if (dot3(d, d) < 0.001f) probe_start = ray_dir; }
  const v2f start_oct = oct_encode(norm3(probe_start)), end_oct = oct_encode(norm3(probe_end));
  const f3 front = oct_center(FM(0.5f, FA(start_oct.x, end_oct.x)), FM(0.5f, FA(start_oct.y, end_oct.y)));
  const float start_depth = FS(encode_oct_depth(dot3(probe_start, front), PR_ZNEAR, PR_ZFAR), 0.0005f);
  const float end_depth = encode_oct_depth(dot3(probe_end, front), PR_ZNEAR, PR_ZFAR);
  const f3 p_start = mk3(start_oct.x, start_oct.y, start_depth), p_end = mk3(end_oct.x, end_oct.y, end_depth);
  const unsigned long long poff = (unsigned long long)probe_index * A.depth.mip[0].ls;
  const f3 p_stop = probe_raymarch(A.depth, poff, p_start, p_end - p_start);
  const float HIT_BIAS = 0.0005f;
  stop_tc = mk2(p_stop.x, p_stop.y);                                                               // valid = (i <= 25): always
  const f3 stop_dir = oct_decode(p_stop.x, p_stop.y);
  const float coef = decode_oct_depth(p_stop.z, PR_ZNEAR, PR_ZFAR) / dot3(stop_dir, front);
  tmin = len3(stop_dir * coef - ray_origin);
  Mip d0 = A.depth.mip[0];
  d0.p += poff;
  const float sampled_depth = sample_r16unorm(d0, p_stop.x, p_stop.y);
  if (p_stop.z > 1) return TRACE_RESULT_MISS;
  if (p_stop.z > FA(sampled_depth, HIT_BIAS)) return TRACE_RESULT_UNKNOWN;
  if (p_stop.z > FS(sampled_depth, HIT_BIAS)) return TRACE_RESULT_HIT;
  return TRACE_RESULT_MISS;
}

// :386-412
__device__ __forceinline__ int trace_probe_one(const ProbeK& k, const ProbeArrays& A, f3 ray_origin, f3 ray_dir, int probe_index, float& tmin, float tmax, v2f& hitexc) {
  const float denom = (float)(k.grid_size - 1u);
  const f3 probe_step = mk3(FS(k.probe_max.x, k.probe_min.x) / denom, FS(k.probe_max.y, k.probe_min.y) / denom, FS(k.probe_max.z, k.probe_min.z) / denom);
  const float gy = (float)((uint32_t)probe_index / k.grid_size), gx = (float)((uint32_t)probe_index % k.grid_size);
  const f3 probeOrigin = k.probe_min + mk3(gx, 0.f, gy) * probe_step;                              // get_probe_position :341-346
  const f3 o = ray_origin - probeOrigin;
  const f3 d = norm3(ray_dir);
  float seg[5];                                                                                     // compute_trace_segments :110-124
  seg[0] = tmin;
  float tx = FM(o.x, -(1.0f / d.x)), ty = FM(o.y, -(1.0f / d.y)), tz = FM(o.z, -(1.0f / d.z));
  swap_min(tx, ty); swap_min(ty, tz); swap_min(tx, ty);
  seg[1] = gclamp(tx, tmin, tmax); seg[2] = gclamp(ty, tmin, tmax); seg[3] = gclamp(tz, tmin, tmax);
  seg[4] = tmax;
#pragma unroll
  for (int i = 0; i < 4; i++)
    if (fabsf(FS(seg[i + 1], seg[i])) >= 0.002f) {
      const int result = trace_segment_hi(A, o, d, seg[i], seg[i + 1], tmin, probe_index, hitexc);
      if (result != TRACE_RESULT_MISS) return result;
    }
  return TRACE_RESULT_MISS;
}

// main() :48-86 and trace() :414-444
__global__ void __launch_bounds__(128) trace_probe_kernel(ProbeK k, LaunchX lx, int cw, int ch, Mip depth, Mip normal, ProbeArrays A, Mip out) {
  const int px = blockIdx.x * 16 + (threadIdx.x & 15), py = VKR_BLOCK_Y * 8 + (threadIdx.x >> 4);
  if (px >= cw || py >= ch) return;
  uint32_t* dst = (uint32_t*)(out.p + (size_t)py * out.pitch) + px;
  const float su = (float)px / (float)cw, sv = (float)py / (float)ch;
  const float pixel_depth = sample_d24(depth, su, sv);
  if (pixel_depth >= 1.f) { *dst = 0u; return; }
  const f3 view_vec = reconstruct_view_vec(su, sv, pixel_depth, k.cam);
  const f3 N = sample_gbuffer_normal(normal, su, sv);
  f3 world_pos = mul_m4(k.inverse_view, view_vec, 1.f);
  world_pos = world_pos + 1e-6f * N;
  const f3 camera_pos = mul_m4(k.inverse_view, mk3(0.f, 0.f, 0.f), 1.f);
  const f3 V = norm3(world_pos - camera_pos);
  world_pos = world_pos - 1e-6f * V;
  const f3 R = reflect3(V, N);

  const float denom = (float)(k.grid_size - 1u);                                                   // get_start_probe_index :333-339
  const float hi = (float)(k.grid_size - 2u);
  const float cx = gclamp(FS(world_pos.x, k.probe_min.x) / (FS(k.probe_max.x, k.probe_min.x) / denom), 0.f, hi);
  const float cz = gclamp(FS(world_pos.z, k.probe_min.z) / (FS(k.probe_max.z, k.probe_min.z) / denom), 0.f, hi);
  const int base_probe = (int)floorf(cz) * (int)k.grid_size + (int)floorf(cx);
  int i = 0, probes_left = 4, hit_probe = -1, result = TRACE_RESULT_UNKNOWN;
  float tmin = 0.f;
  const float tmax = 30.f;
  v2f hit = mk2(0.f, 0.f);
  while (probes_left > 0) {
    const int bx = base_probe % (int)k.grid_size, by = base_probe / (int)k.grid_size;              // next_probe :348-354
    const int probe_id = (by + ((i >> 1) & 1)) * (int)k.grid_size + bx + (i & 1);
    result = trace_probe_one(k, A, world_pos, R, probe_id, tmin, tmax, hit);
    if (result == TRACE_RESULT_UNKNOWN) { i = (i + 3) & 3; probes_left--; }
    else { if (result == TRACE_RESULT_HIT) hit_probe = probe_id; break; }
  }
  uint32_t packed = 0u;
  if (result == TRACE_RESULT_HIT) {
    const float4 c = sample_rgba8unorm(layer_of(A.color, hit_probe), hit.x, hit.y);
    packed = pack_rgba8unorm(c.x, c.y, c.z, c.w);
  }
  *dst = packed;
}

}  // namespace vkr
using namespace vkr;

extern "C" int vkr_cube2oct(const VkrImage* cube_color, uint64_t color_face_stride, const VkrImage* cube_distance, uint64_t distance_face_stride,
                            const VkrImage* oct_color, uint64_t oct_color_layer_stride, const VkrImage* oct_depth, uint64_t oct_depth_layer_stride,
                            uint32_t first_layer, uint32_t probe_count, void* stream) {
  if (!image_ok(cube_color, VKR_FORMAT_R8G8B8A8_SRGB) || !image_ok(cube_distance, VKR_FORMAT_R16_SFLOAT) || !image_ok(oct_color, VKR_FORMAT_R8G8B8A8_UNORM) ||
      !image_ok(oct_depth, VKR_FORMAT_R16_UNORM) || !color_face_stride || !distance_face_stride || !probe_count || probe_count > 65535 ||
      ((first_layer || probe_count > 1) && (!oct_color_layer_stride || !oct_depth_layer_stride))) {
    set_error("cube2oct: null argument, unexpected format or missing stride");
    return VKR_ERR_INVALID_ARGUMENT;
  }
  if (oct_color->width != oct_depth->width || oct_color->height != oct_depth->height || cube_color->width != cube_color->height ||
      cube_distance->width != cube_distance->height) {
    set_error("cube2oct: octahedral colour / depth sizes differ or a cube face is not square");
    return VKR_ERR_SIZE_MISMATCH;
  }
  const int cw = (int)(oct_color->width / 8 * 8), ch = (int)(oct_color->height / 4 * 4);          // probe_renderer.cpp:201
  if (!cw || !ch) return VKR_OK;
  const float* srgb = srgb_decode_table_device();
  if (!srgb) { set_error("cube2oct: sRGB table upload failed"); return VKR_ERR_CUDA; }
  Mip cc = make_mip(cube_color, 0), cd = make_mip(cube_distance, 0), oc = make_mip(oct_color, 0), od = make_mip(oct_depth, 0);
  cc.ls = color_face_stride; cd.ls = distance_face_stride;
  oc.ls = oct_color_layer_stride; od.ls = oct_depth_layer_stride;
  oc.p += (size_t)first_layer * oct_color_layer_stride; od.p += (size_t)first_layer * oct_depth_layer_stride;
  dim3 grid((cw + 31) / 32, (ch + 7) / 8, probe_count);
  cube2oct_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(cw, ch, cc, cd, 6ull * color_face_stride, 6ull * distance_face_stride, oc, od, srgb);
  return check_launch("cube2oct_kernel");
}

extern "C" int vkr_probe_downsample(const VkrImage* probe_depth, uint64_t layer_stride, uint32_t first_layer, uint32_t layer_count, void* stream) {
  if (!image_ok(probe_depth, VKR_FORMAT_R16_UNORM) || !layer_count || layer_count > 65535 || ((first_layer || layer_count > 1) && !layer_stride)) {
    set_error("probe_downsample: null argument, unexpected format or missing stride");
    return VKR_ERR_INVALID_ARGUMENT;
  }
  for (uint32_t i = 1; i < probe_depth->mip_count; i++) {                                          // one task per mip in the reference (:215)
    const int w = (int)(probe_depth->width / (1u << i)), h = (int)(probe_depth->height / (1u << i));   // :226
    if (!w || !h) continue;
    Mip src = make_mip(probe_depth, i - 1), dst = make_mip(probe_depth, i);
    src.ls = dst.ls = layer_stride;
    src.p += (size_t)first_layer * layer_stride; dst.p += (size_t)first_layer * layer_stride;
    dim3 grid((w + 31) / 32, (h + 7) / 8, layer_count);
    probe_downsample_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(src, dst, w, h);
    int rc = check_launch("probe_downsample_kernel");
    if (rc) return rc;
  }
  return VKR_OK;
}

extern "C" int vkr_trace_probe(const VkrProbeTraceParams* p, const VkrImage* depth, const VkrImage* normal, const VkrImage* probe_color,
                               uint64_t color_layer_stride, const VkrImage* probe_depth, uint64_t depth_layer_stride, const VkrImage* out,
                               const VkrLaunchExt* ext, void* stream) {
  if (!p || !image_ok(depth, VKR_FORMAT_D24_UNORM_S8_UINT) || !image_ok(normal, VKR_FORMAT_R16G16_UNORM) || !image_ok(probe_color, VKR_FORMAT_R8G8B8A8_UNORM) ||
      !image_ok(probe_depth, VKR_FORMAT_R16_UNORM) || !image_ok(out, VKR_FORMAT_R8G8B8A8_UNORM) || p->grid_size < 2 || !color_layer_stride || !depth_layer_stride) {
    set_error("trace_probe: null argument, unexpected format, grid_size < 2 or missing stride");
    return VKR_ERR_INVALID_ARGUMENT;
  }
  const int cw = (int)(out->width / 8 * 8), ch = (int)(out->height / 4 * 4);                       // probe_renderer.cpp:381
  if (!cw || !ch) return VKR_OK;
  ExtInfo x;
  int rc = resolve_ext(ext, (uint32_t)ch, 8, &x, "trace_probe");
  if (rc) return rc;
  if (x.layers > 1) { set_error("trace_probe: array layers are not supported"); return VKR_ERR_UNSUPPORTED; }
  ProbeK k;
  k.inverse_view = make_mat(p->inverse_view);
  k.cam = make_cam(p->fovy, p->aspect, p->znear, p->zfar);
  k.probe_min.x = p->probe_min[0]; k.probe_min.y = p->probe_min[1]; k.probe_min.z = p->probe_min[2];
  k.probe_max.x = p->probe_max[0]; k.probe_max.y = p->probe_max[1]; k.probe_max.z = p->probe_max[2];
  k.grid_size = p->grid_size;
  ProbeArrays A;
  A.depth.count = (int)probe_depth->mip_count;
  for (int i = 0; i < VKR_MAX_MIPS; i++) {
    A.depth.mip[i] = (i < A.depth.count) ? make_mip(probe_depth, i) : Mip{nullptr, 0, 0, 0, 0};
    A.depth.mip[i].ls = depth_layer_stride;
  }
  A.color = make_mip(probe_color, 0);
  A.color.ls = color_layer_stride;
  dim3 grid((cw + 15) / 16, x.yblocks, 1);
  trace_probe_kernel<<<grid, 128, 0, (cudaStream_t)stream>>>(k, LaunchX{x.yb0}, cw, ch, make_mip(depth, 0), make_mip(normal, 0), A, make_mip(out, 0));
  return check_launch("trace_probe_kernel");
}
