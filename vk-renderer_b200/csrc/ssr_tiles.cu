// ssr_tiles.cu — the tile path of the stochastic SSR: 8x8 tiles binned into a mirror list and a glossy list by their mean
// roughness, a ray march per listed tile (mirror: 50 cells from the finest level; glossy: 25 cells from one level up), and a
// least-squares plane per tile.  The reference keeps these passes in its tree and leaves them out of AdvancedSSR::run
// (src/advanced_ssr.cpp:547-550); SURVEY §8 f3.
//   "SSSR_Clear"            src/advanced_ssr.cpp:440-452
//   "SSSR_Classification"   src/advanced_ssr.cpp:454-495, shaders/advanced_ssr/classification.comp:36-91
//   "SSSR_trace" (indirect) src/advanced_ssr.cpp:216-302, shaders/advanced_ssr/trace_indirect.comp:44-134
//   "SSSR_Tile_Regression"  src/advanced_ssr.cpp:497-538, shaders/advanced_ssr/regression.comp:22-101
//
// B200 shape: vkCmdDispatchIndirect becomes a persistent grid (a multiple of the SM count) whose CTAs read the tile count from
// device memory and stride over the list — no host read-back between classification and trace, and rays of one list share a
// march depth and a starting level, so a warp's lanes stay together (the point of the binning).  Shared-memory tree sums of
// the shaders keep their association (offset 32 through shared memory, 16..1 as warp shuffles) so the sums match the oracle
// bit for bit.  Everything is unfused: tile membership hangs on `average < threshold`, the march on cell indices.
#include "brdf.cuh"
#include "march.cuh"

namespace vkr {

struct TraceTilesK {
  Mat4 normal_mat;
  CamK cam;
  uint32_t frame_random;
  float max_roughness;
};

struct RegressionK {
  Mat4 camera_to_world;
  CamK cam;
};

// trace_indirect.comp:149-151, bit-defined (see ssr_trace.cu rand_hash)
__device__ __forceinline__ float rand_hash_t(float u, float v) {
  float d = __fadd_rn(__fmul_rn(u, 12.9898f), __fmul_rn(v, 78.233f));
  float s = (float)sin((double)d);
  float p = __fmul_rn(s, 43758.5453f);
  return __fsub_rn(p, floorf(p));
}

// Sum of NV values over the 64 threads of a tile in the shaders' tree order: g[t] += g[t + offset], offset = 32, 16 .. 1.
// t = thread index in the tile (two warps); sh holds NV x 32 floats per tile.  The total is valid in t == 0.
template <int NV>
__device__ __forceinline__ void tile_tree_sum(float (&v)[NV], float* sh, int t) {
  if (t >= 32) {
#pragma unroll
    for (int i = 0; i < NV; i++) sh[i * 32 + (t - 32)] = v[i];
  }
  __syncthreads();
  if (t < 32) {
#pragma unroll
    for (int i = 0; i < NV; i++) v[i] = FA(v[i], sh[i * 32 + t]);
  }
  __syncthreads();
#pragma unroll
  for (int off = 16; off > 0; off >>= 1) {
#pragma unroll
    for (int i = 0; i < NV; i++) v[i] = FA(v[i], __shfl_down_sync(0xFFFFFFFFu, v[i], off));
  }
}

__global__ void sssr_clear_indirect_kernel(uint32_t* a, uint32_t* b) {
  if (threadIdx.x < 3) {
    const uint32_t v = threadIdx.x == 0 ? 0u : 1u;      // VkDispatchIndirectCommand {0, 1, 1}
    a[threadIdx.x] = v;
    b[threadIdx.x] = v;
  }
}

// One tile per 64 threads, four tiles (32 x 8 pixels) per CTA.
__global__ void __launch_bounds__(256) sssr_classification_kernel(int tex_w, int tex_h, float max_roughness, float glossy_value, Mip material,
                                                                  const float* __restrict__ srgb, int* __restrict__ reflective_tiles,
                                                                  int* __restrict__ glossy_tiles, uint32_t* reflective_count, uint32_t* glossy_count) {
  __shared__ float sh[4][32];
  const int slot = threadIdx.x >> 6, t = threadIdx.x & 63;
  const int tile_x = blockIdx.x * 4 + slot, tile_y = blockIdx.y;
  const int px = tile_x * 8 + (t & 7), py = tile_y * 8 + (t >> 3);
  float v[1];
  v[0] = 1.f;
  if (px < tex_w && py < tex_h) v[0] = sample_srgb8(material, (float)px / (float)tex_w, (float)py / (float)tex_h, srgb).y;    // :39-46
  v[0] = gmix(0.f, max_roughness, v[0]);                                                                                       // :49
  tile_tree_sum<1>(v, sh[slot], t);
  const int tiles_x = (tex_w + 7) / 8;
  if (t == 0 && tile_x < tiles_x) {                                                                                            // :78-90
    const float average_roughness = v[0] / 64.0f;
    const int tile_index = tile_y * tiles_x + tile_x;
    if (average_roughness < glossy_value) reflective_tiles[atomicAdd(reflective_count, 1u)] = tile_index;
    else glossy_tiles[atomicAdd(glossy_count, 1u)] = tile_index;
  }
}

// MIP = most detailed view level the march may reach (0 mirror, 1 glossy); MAX_IT = cell budget.  Two tiles per CTA pass.
template <int MIP, int MAX_IT, bool THICKNESS_TEST>
__global__ void __launch_bounds__(128) sssr_trace_indirect_kernel(TraceTilesK k, PyramidView P, Mip normal, Mip material, const float4* __restrict__ halton,
                                                                  const int* __restrict__ g_tiles, const uint32_t* __restrict__ indirect, uint32_t max_tiles,
                                                                  Mip out_rays, const float* __restrict__ srgb) {
  uint32_t count = __ldg(indirect);
  if (count > max_tiles) count = max_tiles;
  const int slot = threadIdx.x >> 6, t = threadIdx.x & 63;
  const int tile_width = (out_rays.w + 7) / 8;                                                         // :48
  const float tsx = (float)out_rays.w, tsy = (float)out_rays.h;
  const CamK cam = k.cam;
  const Mip d0 = P.mip[0];
  for (uint32_t wg = blockIdx.x * 2 + slot; wg < count; wg += gridDim.x * 2) {
    const int tile_index = __ldg(g_tiles + wg);
    const int px = 8 * (tile_index % tile_width) + (t & 7), py = 8 * (tile_index / tile_width) + (t >> 3);
    if (px >= out_rays.w || py >= out_rays.h || px < 0 || py < 0) continue;
    const float su = (float)px / tsx, sv = (float)py / tsy;                                            // :53

    float roughness = sample_srgb8(material, su, sv, srgb).y;                                          // :59-63
    roughness = gmix(0.0f, k.max_roughness, roughness);
    roughness = FM(roughness, roughness);

    const float pixel_depth = sample_d24(d0, su, sv);
    const f3 pixel_normal = norm3(mul_m4(k.normal_mat, sample_gbuffer_normal(normal, su, sv), 0.f));
    const f3 view_vec = reconstruct_view_vec(su, sv, pixel_depth, cam);

    const uint32_t base_index = (uint32_t)(rand_hash_t(su, sv) * 128.0f);                              // :71-73
    const uint32_t index = (base_index + k.frame_random) & 127u;
    const float4 hv = __ldg(halton + index);
    f3 tangent = get_tangent(pixel_normal);
    const f3 bitangent = norm3(cross3(pixel_normal, tangent));
    tangent = norm3(cross3(bitangent, pixel_normal));

    const f3 w0 = -norm3(view_vec);
    const f3 view_dir = mk3(dot3(w0, tangent), dot3(w0, bitangent), dot3(w0, pixel_normal));
    const f3 brdf_norm = sampleGGXVNDF(view_dir, roughness, roughness, hv.x, hv.y);
    const f3 N = brdf_norm.x * tangent + brdf_norm.y * bitangent + brdf_norm.z * pixel_normal;
    const f3 R = reflect3(view_vec, N);

    f3 ray_start = project_view_vec(view_vec + 0.001f * pixel_normal, cam);                            // :89-94
    ray_start.z = FS(ray_start.z, 0.0001f);
    f3 ray_dir = project_view_vec(view_vec + R, cam) - ray_start;
    ray_dir = ray_dir * (FS(1.f, ray_start.z) / ray_dir.z);

    // ---- hierarchical_raymarch(DEPTH, ray_start, ray_dir, MIP, MAX_IT), screen_trace.glsl:51-100 ----
    const f3 origin = ray_start, direction = ray_dir;
    const f3 inv_direction = mk3(direction.x != 0 ? 1.0f / direction.x : MAX_T_FLOAT, direction.y != 0 ? 1.0f / direction.y : MAX_T_FLOAT,
                                 direction.z != 0 ? 1.0f / direction.z : MAX_T_FLOAT);
    int current_mip = MIP;
    const float ssx = (float)d0.w, ssy = (float)d0.h;
    const float mip_scale = MIP ? 0.5f : 1.0f;                  // pow(0.5, most_detailed_mip)
    float resx = FM(ssx, mip_scale), resy = FM(ssy, mip_scale);
    float rinvx = 1.0f / resx, rinvy = 1.0f / resy;
    float uox = FM(0.005f, MIP ? 2.0f : 1.0f) / ssx, uoy = FM(0.005f, MIP ? 2.0f : 1.0f) / ssy;
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
    while (i < MAX_IT && current_mip >= MIP) {
      float mpx = FM(resx, position.x), mpy = FM(resy, position.y);
      float surface_z = fetch_pyramid(P, 0ull, (int)mpx, (int)mpy, current_mip);
      bool skipped = advance_ray(origin, direction, inv_direction, mpx, mpy, rinvx, rinvy, fox, foy, uox, uoy, surface_z, position, current_t);
      current_mip += skipped ? 1 : -1;
      float f = skipped ? 0.5f : 2.0f, g = skipped ? 2.0f : 0.5f;
      resx *= f; resy *= f; rinvx *= g; rinvy *= g;
      ++i;
    }
    const f3 out_ray = position;
    bool valid_hit = true;                                      // (i <= max) always holds, App. A.6 #8

    {                                                                                                   // :107-112
      float sx = FM(fabsf(FS(out_ray.x, ray_start.x)), tsx), sy = FM(fabsf(FS(out_ray.y, ray_start.y)), tsy);
      if (gmax(sx, sy) < 2.0f) valid_hit = false;
    }
    if (valid_hit) {                                                                                    // :114-122
      f3 hit_normal = mul_m4(k.normal_mat, sample_gbuffer_normal(normal, out_ray.x, out_ray.y), 0.f);
      if (dot3(hit_normal, R) > 0 || dot3(pixel_normal, R) < 0) valid_hit = false;
    }
    if (THICKNESS_TEST && valid_hit) {                                                                  // :124-131
      float hit_depth = sample_d24(d0, out_ray.x, out_ray.y);
      float hit_z = linearize_depth2(hit_depth, cam.znear, cam.zfar);
      float ray_z = linearize_depth2(out_ray.z, cam.znear, cam.zfar);
      if (ray_z > FA(hit_z, 0.3f) || ray_z < FS(hit_z, 0.1f)) valid_hit = false;
    }
    uint32_t lo = to_unorm(out_ray.x, 65535.f) | (to_unorm(out_ray.y, 65535.f) << 16);                  // :133-134
    uint32_t hi = to_unorm(out_ray.z, 65535.f) | (to_unorm(valid_hit ? pixel_depth : 1.0f, 65535.f) << 16);
    *((uint2*)(out_rays.p + (size_t)py * out_rays.pitch) + px) = make_uint2(lo, hi);
  }
}

// One tile per 64-thread CTA.
__global__ void __launch_bounds__(64) sssr_tile_regression_kernel(RegressionK k, Mip depth, Mip out_planes) {
  __shared__ float sh[9 * 32];
  __shared__ float plane_sh[3];
  const int t = threadIdx.x;
  const int px = blockIdx.x * 8 + (t & 7), py = blockIdx.y * 8 + (t >> 3);
  f3 view_vec = mk3(0.f, 0.f, 0.f);
  if (px < depth.w && py < depth.h) {                                                                   // :28-37
    const float pixel_depth = fetch_d24(depth, px, py);
    view_vec = reconstruct_view_vec((float)px / (float)depth.w, (float)py / (float)depth.h, pixel_depth, k.cam);
    const f3 world_vec = mul_m4(k.camera_to_world, view_vec, 1.f);
    const f3 world_origin = mul_m4(k.camera_to_world, mk3(0.f, 0.f, 0.f), 1.f);
    view_vec = world_vec - world_origin;
  }
  float v[9] = {view_vec.x, view_vec.y, view_vec.z,                                                     // :40-44
                FM(view_vec.x, view_vec.x), FM(view_vec.y, view_vec.y), FM(view_vec.z, view_vec.z),
                FM(view_vec.x, view_vec.y), FM(view_vec.x, view_vec.z), FM(view_vec.y, view_vec.z)};
  tile_tree_sum<9>(v, sh, t);
  if (t == 0) {                                                                                         // :64-73, inverse(mat3) = adjugate / det
    const float a = v[3], b = v[6], c = v[7], d = v[4], e = v[8], f = v[5];
    const float c00 = FS(FM(d, f), FM(e, e)), c01 = FS(FM(c, e), FM(b, f)), c02 = FS(FM(b, e), FM(c, d));
    const float c11 = FS(FM(a, f), FM(c, c)), c12 = FS(FM(b, c), FM(a, e)), c22 = FS(FM(a, d), FM(b, b));
    const float det = FA(FA(FM(a, c00), FM(b, c01)), FM(c, c02));
    const float inv_det = 1.0f / det;
    plane_sh[0] = FA(FA(FM(FM(c00, inv_det), v[0]), FM(FM(c01, inv_det), v[1])), FM(FM(c02, inv_det), v[2]));
    plane_sh[1] = FA(FA(FM(FM(c01, inv_det), v[0]), FM(FM(c11, inv_det), v[1])), FM(FM(c12, inv_det), v[2]));
    plane_sh[2] = FA(FA(FM(FM(c02, inv_det), v[0]), FM(FM(c12, inv_det), v[1])), FM(FM(c22, inv_det), v[2]));
  }
  __syncthreads();
  const f3 plane = mk3(plane_sh[0], plane_sh[1], plane_sh[2]);
  float s[1];
  s[0] = FS(dot3(plane, view_vec), 1.f);                                                                // :76-82
  s[0] = FM(s[0], s[0]);
  if (isnan(s[0])) s[0] = 1e10f;
  tile_tree_sum<1>(s, sh, t);
  if (t == 0 && (int)blockIdx.x < out_planes.w && (int)blockIdx.y < out_planes.h)                       // :98-101
    *((float4*)(out_planes.p + (size_t)blockIdx.y * out_planes.pitch) + blockIdx.x) = make_float4(plane.x, plane.y, plane.z, s[0] / 64.0f);
}

}  // namespace vkr
using namespace vkr;

static int sm_count() {
  int dev = 0, n = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0) return 0;
  return n;
}

extern "C" int vkr_sssr_clear_indirect(void* reflective_indirect, void* glossy_indirect, void* stream) {
  if (!reflective_indirect || !glossy_indirect) { set_error("sssr_clear_indirect: null argument"); return VKR_ERR_INVALID_ARGUMENT; }
  sssr_clear_indirect_kernel<<<1, 32, 0, (cudaStream_t)stream>>>((uint32_t*)reflective_indirect, (uint32_t*)glossy_indirect);
  return check_launch("sssr_clear_indirect_kernel");
}

extern "C" int vkr_sssr_classification(const VkrClassifyParams* p, const VkrImage* material, int32_t* reflective_tiles, int32_t* glossy_tiles,
                                       void* reflective_indirect, void* glossy_indirect, void* stream) {
  if (!p || !reflective_tiles || !glossy_tiles || !reflective_indirect || !glossy_indirect || !image_ok(material, VKR_FORMAT_R8G8B8A8_SRGB) ||
      p->width <= 0 || p->height <= 0) {
    set_error("sssr_classification: null argument, empty extent or unexpected format");
    return VKR_ERR_INVALID_ARGUMENT;
  }
  const float* srgb = srgb_decode_table_device();
  if (!srgb) { set_error("sssr_classification: sRGB table upload failed"); return VKR_ERR_CUDA; }
  const int tiles_x = (p->width + 7) / 8, tiles_y = (p->height + 7) / 8;
  dim3 grid((tiles_x + 3) / 4, tiles_y);
  sssr_classification_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(p->width, p->height, p->max_roughness, p->glossy_value, make_mip(material, 0), srgb,
                                                                      reflective_tiles, glossy_tiles, (uint32_t*)reflective_indirect, (uint32_t*)glossy_indirect);
  return check_launch("sssr_classification_kernel");
}

extern "C" int vkr_sssr_trace_indirect(const VkrTraceParams* p, uint32_t reflection_type, const VkrImage* depth, const VkrImage* dnormal,
                                       const VkrImage* material, const float* halton_vec4x128, const int32_t* tiles, const void* indirect,
                                       uint32_t max_tiles, const VkrImage* out_rays, void* stream) {
  if (!p || !halton_vec4x128 || !tiles || !indirect || reflection_type > 1 || !image_ok(depth, VKR_FORMAT_D24_UNORM_S8_UINT) || depth->mip_count < 2 ||
      !image_ok(dnormal, VKR_FORMAT_R16G16_UNORM) || !image_ok(material, VKR_FORMAT_R8G8B8A8_SRGB) || !image_ok(out_rays, VKR_FORMAT_R16G16B16A16_UNORM)) {
    set_error("sssr_trace_indirect: null argument, reflection_type > 1 or unexpected format");
    return VKR_ERR_INVALID_ARGUMENT;
  }
  if (max_tiles == 0) return VKR_OK;
  const float* srgb = srgb_decode_table_device();
  if (!srgb) { set_error("sssr_trace_indirect: sRGB table upload failed"); return VKR_ERR_CUDA; }
  const int sms = sm_count();
  if (!sms) { set_error("sssr_trace_indirect: no CUDA device"); return VKR_ERR_CUDA; }
  TraceTilesK k;
  k.normal_mat = make_mat(p->normal_mat);
  k.cam = make_cam(p->fovy, p->aspect, p->znear, p->zfar);
  k.frame_random = p->frame_random; k.max_roughness = p->max_roughness;
  PyramidView P;
  P.count = (int)depth->mip_count - 1;                     // view = mips 1..L-1 (advanced_ssr.cpp:254)
  for (int i = 0; i < VKR_MAX_MIPS; i++) P.mip[i] = (i < P.count) ? make_mip(depth, i + 1) : Mip{nullptr, 0, 0, 0, 0};
  // persistent grid: 8 CTAs of 128 threads per SM, never more CTAs than tile pairs
  uint32_t ctas = (uint32_t)sms * 8u, need = (max_tiles + 1u) / 2u;
  if (ctas > need) ctas = need;
  cudaStream_t st = (cudaStream_t)stream;
  if (reflection_type == 0)
    sssr_trace_indirect_kernel<0, 50, true><<<ctas, 128, 0, st>>>(k, P, make_mip(dnormal, 0), make_mip(material, 0), (const float4*)halton_vec4x128, tiles,
                                                                  (const uint32_t*)indirect, max_tiles, make_mip(out_rays, 0), srgb);
  else
    sssr_trace_indirect_kernel<1, 25, false><<<ctas, 128, 0, st>>>(k, P, make_mip(dnormal, 0), make_mip(material, 0), (const float4*)halton_vec4x128, tiles,
                                                                   (const uint32_t*)indirect, max_tiles, make_mip(out_rays, 0), srgb);
  return check_launch("sssr_trace_indirect_kernel");
}

extern "C" int vkr_sssr_tile_regression(const VkrRegressionParams* p, const VkrImage* depth, const VkrImage* out_planes, void* stream) {
  if (!p || !image_ok(depth, VKR_FORMAT_D24_UNORM_S8_UINT) || depth->mip_count < 2 || !image_ok(out_planes, VKR_FORMAT_R32G32B32A32_SFLOAT)) {
    set_error("sssr_tile_regression: null argument or unexpected format");
    return VKR_ERR_INVALID_ARGUMENT;
  }
  RegressionK k;
  k.camera_to_world = make_mat(p->camera_to_world);
  k.cam = make_cam(p->fovy, p->aspect, p->znear, p->zfar);
  const Mip d1 = make_mip(depth, 1);                        // view = mip 1 only (advanced_ssr.cpp:524)
  dim3 grid((d1.w + 7) / 8, (d1.h + 7) / 8);
  sssr_tile_regression_kernel<<<grid, 64, 0, (cudaStream_t)stream>>>(k, d1, make_mip(out_planes, 0));
  return check_launch("sssr_tile_regression_kernel");
}

// ================================================================================================
// vkr_sssr_trace_tiles_static — the whole tile path of the trace (SSSR_Clear -> SSSR_Classification -> the two indirect
// SSSR_trace dispatches) in ONE call that allocates nothing.  No counterpart in the reference; added
// for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses
// that forbid it).
//
// The scratch the separate entries make the caller allocate — two tile lists, two indirect commands — and the sRGB decode
// table (which the other entries cudaMalloc lazily, api.cu srgb_decode_table_device) live in ONE file-scope device array of
// fixed size.  Every GPU that loads the library holds it (VKR_SSSR_STATIC_SCRATCH_BYTES, about 513 KB) until exit.
//   words [0, 4)                      reflective indirect command {count, 1, 1, -}
//   words [4, 8)                      glossy indirect command
//   words [64, 64 + T)                reflective tile list     (T = VKR_SSSR_STATIC_MAX_TILES)
//   words [64 + T, 64 + 2T)           glossy tile list
//   words [64 + 2T, 64 + 2T + 256)    sRGB8 -> linear table (floats), written once per device at set-up
// In HOST code the array's name is a host-side stand-in, not a device address: the address the kernels get comes from
// cudaGetSymbolAddress, once per device with that device current (static_scratch()).
// Every call shares the array, which keeps the last call's contents: the two counters are reset by the clear kernel at the
// start of every call, list entries beyond the counters are never read.  Calls that OVERLAP on the GPU (two streams, or two
// host threads on one device) race on it: issue them on one stream or order them with events.
#include <mutex>

namespace vkr {

#define SSSR_STATIC_LIST_WORD0 64u
#define SSSR_STATIC_WORDS (SSSR_STATIC_LIST_WORD0 + 2u * VKR_SSSR_STATIC_MAX_TILES + 256u)
static_assert(SSSR_STATIC_WORDS * 4u == VKR_SSSR_STATIC_SCRATCH_BYTES, "header and layout disagree");

static __device__ __align__(256) uint32_t g_sssr_static_scratch[SSSR_STATIC_WORDS];

static std::mutex g_sssr_static_mutex;
static uint32_t* g_sssr_static_ptr[64] = {nullptr};     // device address of the array on each device, filled at set-up

// Set-up, once per device, with that device current: the array's device address + the sRGB table (the same bits as
// srgb_decode_table_device builds).  No allocation: cudaGetSymbolAddress looks the symbol up, cudaMemcpy writes into it.
static uint32_t* static_scratch() {
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) { set_error("sssr_trace_tiles_static: no current CUDA device"); return nullptr; }
  std::lock_guard<std::mutex> lock(g_sssr_static_mutex);
  if (g_sssr_static_ptr[dev]) return g_sssr_static_ptr[dev];
  uint32_t* ptr = nullptr;
  cudaError_t e = cudaGetSymbolAddress((void**)&ptr, g_sssr_static_scratch);
  if (e != cudaSuccess || !ptr) { set_error("sssr_trace_tiles_static: cudaGetSymbolAddress: %s", cudaGetErrorString(e)); return nullptr; }
  float host[256];
  for (int i = 0; i < 256; i++) {
    const double c = i / 255.0;
    host[i] = (float)((c <= 0.04045) ? c / 12.92 : std::pow((c + 0.055) / 1.055, 2.4));
  }
  e = cudaMemcpy(ptr + SSSR_STATIC_LIST_WORD0 + 2u * VKR_SSSR_STATIC_MAX_TILES, host, sizeof(host), cudaMemcpyHostToDevice);
  if (e != cudaSuccess) { set_error("sssr_trace_tiles_static: sRGB table upload: %s", cudaGetErrorString(e)); return nullptr; }
  g_sssr_static_ptr[dev] = ptr;
  return ptr;
}

}  // namespace vkr

extern "C" int vkr_sssr_trace_tiles_static(const VkrTraceParams* p, float glossy_value, const VkrImage* depth, const VkrImage* dnormal,
                                           const VkrImage* material, const float* halton_vec4x128, const VkrImage* out_rays, void* stream) {
  if (!p || !halton_vec4x128 || !image_ok(depth, VKR_FORMAT_D24_UNORM_S8_UINT) || depth->mip_count < 2 || !image_ok(dnormal, VKR_FORMAT_R16G16_UNORM) ||
      !image_ok(material, VKR_FORMAT_R8G8B8A8_SRGB) || !image_ok(out_rays, VKR_FORMAT_R16G16B16A16_UNORM)) {
    set_error("sssr_trace_tiles_static: null argument or unexpected format");
    return VKR_ERR_INVALID_ARGUMENT;
  }
  const int w = (int)out_rays->width, h = (int)out_rays->height;
  const int tiles_x = (w + 7) / 8, tiles_y = (h + 7) / 8;
  const unsigned long long n_tiles = (unsigned long long)tiles_x * (unsigned long long)tiles_y;
  if (n_tiles > VKR_SSSR_STATIC_MAX_TILES) {
    set_error("sssr_trace_tiles_static: %dx%d rays = %llu tiles of 8x8, beyond the static scratch's maximum of %u tiles (VKR_SSSR_STATIC_MAX_TILES); "
              "use vkr_sssr_classification / vkr_sssr_trace_indirect with caller-allocated lists", w, h, n_tiles, (unsigned)VKR_SSSR_STATIC_MAX_TILES);
    return VKR_ERR_UNSUPPORTED;
  }
  uint32_t* S = static_scratch();
  if (!S) return VKR_ERR_CUDA;
  const int sms = sm_count();
  if (!sms) { set_error("sssr_trace_tiles_static: no CUDA device"); return VKR_ERR_CUDA; }
  uint32_t* refl_cmd = S;
  uint32_t* glossy_cmd = S + 4;
  int* refl_tiles = (int*)(S + SSSR_STATIC_LIST_WORD0);
  int* glossy_tiles = refl_tiles + VKR_SSSR_STATIC_MAX_TILES;
  const float* srgb = (const float*)(S + SSSR_STATIC_LIST_WORD0 + 2u * VKR_SSSR_STATIC_MAX_TILES);
  cudaStream_t st = (cudaStream_t)stream;
  const uint32_t max_tiles = (uint32_t)n_tiles;

  sssr_clear_indirect_kernel<<<1, 32, 0, st>>>(refl_cmd, glossy_cmd);                       // the counters must start at zero: every call
  dim3 cgrid((tiles_x + 3) / 4, tiles_y);
  sssr_classification_kernel<<<cgrid, 256, 0, st>>>(w, h, p->max_roughness, glossy_value, make_mip(material, 0), srgb, refl_tiles, glossy_tiles, refl_cmd, glossy_cmd);
  TraceTilesK k;
  k.normal_mat = make_mat(p->normal_mat);
  k.cam = make_cam(p->fovy, p->aspect, p->znear, p->zfar);
  k.frame_random = p->frame_random; k.max_roughness = p->max_roughness;
  PyramidView P;
  P.count = (int)depth->mip_count - 1;
  for (int i = 0; i < VKR_MAX_MIPS; i++) P.mip[i] = (i < P.count) ? make_mip(depth, i + 1) : Mip{nullptr, 0, 0, 0, 0};
  uint32_t ctas = (uint32_t)sms * 8u, need = (max_tiles + 1u) / 2u;
  if (ctas > need) ctas = need;
  sssr_trace_indirect_kernel<0, 50, true><<<ctas, 128, 0, st>>>(k, P, make_mip(dnormal, 0), make_mip(material, 0), (const float4*)halton_vec4x128, refl_tiles,
                                                                refl_cmd, max_tiles, make_mip(out_rays, 0), srgb);
  sssr_trace_indirect_kernel<1, 25, false><<<ctas, 128, 0, st>>>(k, P, make_mip(dnormal, 0), make_mip(material, 0), (const float4*)halton_vec4x128, glossy_tiles,
                                                                 glossy_cmd, max_tiles, make_mip(out_rays, 0), srgb);
  return check_launch("sssr_trace_tiles_static");
}
