// hiz.cu — DownsamplePass: half-res G-buffer + min-depth Hi-Z pyramid on 24-bit integers.
//
// Replaces the tasks "DownsampleGbuffer" and "DownsampleDepth" x (L-2)
// (reference src/downsample_pass.cpp:25-131; shaders/advanced_ssr/downsample_gbuffer.frag:12-37,
// shaders/advanced_ssr/depth_mips.frag:7-15).  The reference issues L-1 barrier-separated
// fullscreen draws; here one launch builds mips 1..6 of every 64x64 tile in registers /
// warp shuffles / shared memory and a one-CTA tail launch finishes the small top levels.
//
// Integer semantics: the shader's float(k)/(2^24-1) compare is
// order-isomorphic to comparing k; gl_FragDepth re-quantisation is the identity, so the
// pyramid is the exact min of the 24-bit codes.  Mip i texel (x,y) covers the full-res block
// [x*2^i,(x+1)*2^i) x [y*2^i,(y+1)*2^i), which always lies inside the image (floor sizes),
// except when a source level has a dimension of 1: the +1 fetch is out of range (= 0).
#include "common.cuh"

namespace vkr {

struct HizLevels {
  Mip mip[VKR_MAX_MIPS];
  int count;
};

#define HIZ_ABSENT 0xFFFFFFFFu

// level i is "degenerate" when its source level i-1 has a dimension of 1.
__device__ __forceinline__ bool degenerate(const HizLevels& L, int i) { return L.mip[i - 1].w == 1 || L.mip[i - 1].h == 1; }

// loff: byte offset of this launch's array layer inside the depth allocation (0 for a single image)
__device__ __forceinline__ void st_level(const HizLevels& L, size_t loff, int i, int x, int y, uint32_t v) {
  if (i < L.count && x < L.mip[i].w && y < L.mip[i].h) {
    if (degenerate(L, i)) v = 0u;
    *((uint32_t*)(L.mip[i].p + loff + (size_t)y * L.mip[i].pitch) + x) = v;
  }
}

// 256 threads, one 64x64 mip-0 tile per CTA.  Thread (tx,ty) of a 16x16 grid owns a 4x4
// texel block; a warp owns an 8x4 patch of threads (32x16 texels) so mips 3 and 4 are
// shuffle reductions and only mips 5/6 go through shared memory.
template <bool PAYLOAD>
__global__ void __launch_bounds__(256) hiz_tile_kernel(HizLevels L, int src, int levels, Mip normal, Mip velocity, Mip out_normal, Mip out_velocity) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int tx = (warp & 1) * 8 + (lane & 7), ty = (warp >> 1) * 4 + (lane >> 3);
  const size_t loff = (size_t)blockIdx.z * L.mip[0].ls;
  if (PAYLOAD) { normal = layer_of(normal, blockIdx.z); velocity = layer_of(velocity, blockIdx.z); out_normal = layer_of(out_normal, blockIdx.z); out_velocity = layer_of(out_velocity, blockIdx.z); }
  Mip s = L.mip[src];
  s.p += loff;
  const int x0 = blockIdx.x * 64 + tx * 4, y0 = blockIdx.y * 64 + ty * 4;

  // 4 rows x uint4 of source depth; absent texels are neutral for min.
  uint32_t d[4][4];
#pragma unroll
  for (int r = 0; r < 4; r++) {
    uint4 v = make_uint4(HIZ_ABSENT, HIZ_ABSENT, HIZ_ABSENT, HIZ_ABSENT);
    if (y0 + r < s.h && (unsigned)(x0 * 4) < s.pitch) v = __ldg((const uint4*)(s.p + (size_t)(y0 + r) * s.pitch) + (x0 >> 2));
    d[r][0] = (x0 + 0 < s.w) ? (v.x & 0xFFFFFFu) : HIZ_ABSENT;
    d[r][1] = (x0 + 1 < s.w) ? (v.y & 0xFFFFFFu) : HIZ_ABSENT;
    d[r][2] = (x0 + 2 < s.w) ? (v.z & 0xFFFFFFu) : HIZ_ABSENT;
    d[r][3] = (x0 + 3 < s.w) ? (v.w & 0xFFFFFFu) : HIZ_ABSENT;
  }

  // level src+1: 2x2 texels per thread (+ payload select, downsample_gbuffer.frag:24-36)
  uint32_t m1[2][2];
  Mip l1 = L.mip[src + 1];
  l1.p += loff;
#pragma unroll
  for (int j = 0; j < 2; j++)
#pragma unroll
    for (int i = 0; i < 2; i++) {
      uint32_t a = d[2 * j][2 * i], b = d[2 * j][2 * i + 1], c = d[2 * j + 1][2 * i], e = d[2 * j + 1][2 * i + 1];
      uint32_t m = min(min(a, b), min(c, e));
      m1[j][i] = m;
      if (PAYLOAD) {
        int ox = 0, oy = 0;
        if (m == b) { ox = 1; oy = 0; }
        else if (m == c) { ox = 0; oy = 1; }
        else if (m == e) { ox = 1; oy = 1; }
        int X = (x0 >> 1) + i, Y = (y0 >> 1) + j;
        if (X < l1.w && Y < l1.h) {
          int sx = x0 + 2 * i + ox, sy = y0 + 2 * j + oy;
          uint32_t nb = (sx < normal.w && sy < normal.h) ? ld32(normal, sx, sy) : 0u;
          uint32_t vb = (sx < velocity.w && sy < velocity.h) ? ld32(velocity, sx, sy) : 0u;
          *((uint32_t*)(out_normal.p + (size_t)Y * out_normal.pitch) + X) = nb;
          *((uint32_t*)(out_velocity.p + (size_t)Y * out_velocity.pitch) + X) = vb;
        }
      }
    }
  {
    int X = x0 >> 1, Y = y0 >> 1;
#pragma unroll
    for (int j = 0; j < 2; j++) {
      if (Y + j < l1.h && X + 1 < l1.w && !degenerate(L, src + 1)) {
        *((uint2*)(l1.p + (size_t)(Y + j) * l1.pitch + (size_t)X * 4)) = make_uint2(m1[j][0], m1[j][1]);
      } else {
        st_level(L, loff, src + 1, X, Y + j, m1[j][0]);
        st_level(L, loff, src + 1, X + 1, Y + j, m1[j][1]);
      }
    }
  }
  if (levels < 2) return;
  uint32_t m = min(min(m1[0][0], m1[0][1]), min(m1[1][0], m1[1][1]));
  st_level(L, loff, src + 2, x0 >> 2, y0 >> 2, m);
  if (levels < 3) return;
  m = min(m, __shfl_xor_sync(0xFFFFFFFFu, m, 1));
  m = min(m, __shfl_xor_sync(0xFFFFFFFFu, m, 8));
  if ((lane & 9) == 0) st_level(L, loff, src + 3, x0 >> 3, y0 >> 3, m);
  if (levels < 4) return;
  m = min(m, __shfl_xor_sync(0xFFFFFFFFu, m, 2));
  m = min(m, __shfl_xor_sync(0xFFFFFFFFu, m, 16));
  if ((lane & 27) == 0) st_level(L, loff, src + 4, x0 >> 4, y0 >> 4, m);
  if (levels < 5) return;
  // warp = 2x1 level-4 texels (lanes 0 and 4); level-5 texel = 2 (x) x 2 (y) of them.
  __shared__ uint32_t s4[4][4];  // [level-4 row][level-4 column] of the tile
  if ((lane & 27) == 0) s4[warp >> 1][(warp & 1) * 2 + (lane >> 2)] = m;
  __syncthreads();
  if (threadIdx.x < 4) {
    int cx = threadIdx.x & 1, cy = threadIdx.x >> 1;
    uint32_t v = min(min(s4[2 * cy][2 * cx], s4[2 * cy][2 * cx + 1]), min(s4[2 * cy + 1][2 * cx], s4[2 * cy + 1][2 * cx + 1]));
    st_level(L, loff, src + 5, blockIdx.x * 2 + cx, blockIdx.y * 2 + cy, v);
    if (levels >= 6) {
      v = min(v, __shfl_xor_sync(0xFu, v, 1));
      v = min(v, __shfl_xor_sync(0xFu, v, 2));
      if (threadIdx.x == 0) st_level(L, loff, src + 6, blockIdx.x, blockIdx.y, v);
    }
  }
}

// One CTA walks the remaining (small) levels; plain loads so the CTA sees its own stores.
__global__ void __launch_bounds__(1024) hiz_tail_kernel(HizLevels L, int first) {
  const size_t loff = (size_t)blockIdx.x * L.mip[0].ls;     // one CTA per array layer
  for (int i = first; i < L.count; i++) {
    Mip s = L.mip[i - 1], o = L.mip[i];
    s.p += loff; o.p += loff;
    const int n = o.w * o.h;
    for (int t = threadIdx.x; t < n; t += blockDim.x) {
      int x = t % o.w, y = t / o.w;
      uint32_t v = 0xFFFFFFFFu;
#pragma unroll
      for (int k = 0; k < 4; k++) {
        int sx = 2 * x + (k & 1), sy = 2 * y + (k >> 1);
        uint32_t e = (sx < s.w && sy < s.h) ? (*((const volatile uint32_t*)(s.p + (size_t)sy * s.pitch) + sx) & 0xFFFFFFu) : 0u;
        v = min(v, e);
      }
      *((uint32_t*)(o.p + (size_t)y * o.pitch) + x) = v;
    }
    __syncthreads();
  }
}

static HizLevels make_levels(const VkrImage* depth, uint64_t layer_stride) {
  HizLevels L;
  L.count = (int)depth->mip_count;
  for (int i = 0; i < VKR_MAX_MIPS; i++) {
    L.mip[i] = (i < L.count) ? make_mip(depth, i) : Mip{nullptr, 0, 0, 0, 0};
    L.mip[i].ls = layer_stride;
  }
  return L;
}

// Builds levels src+1 .. L-1.  payload images may be null (depth only).
static int hiz_run(const VkrImage* depth, int src, const VkrImage* normal, const VkrImage* velocity,
                   const VkrImage* out_normal, const VkrImage* out_velocity, int max_levels, cudaStream_t st, const ExtInfo& x = ExtInfo()) {
  HizLevels L = make_levels(depth, x.stride(0));
  int cur = src;
  bool payload = normal != nullptr;
  const int TAIL_TEXELS = 16384;
  while (cur < L.count - 1) {
    int remaining = L.count - 1 - cur;
    if (max_levels > 0 && remaining > max_levels) remaining = max_levels;
    if (!payload && L.mip[cur].w * L.mip[cur].h <= TAIL_TEXELS && max_levels <= 0) {
      hiz_tail_kernel<<<x.layers, 1024, 0, st>>>(L, cur + 1);
      return check_launch("hiz_tail_kernel");
    }
    int levels = remaining < 6 ? remaining : 6;
    dim3 grid((L.mip[cur].w + 63) / 64, (L.mip[cur].h + 63) / 64, x.layers);
    Mip none{nullptr, 0, 0, 0, 0};
    if (payload)
      hiz_tile_kernel<true><<<grid, 256, 0, st>>>(L, cur, levels, make_mip(normal, 0, x, 1), make_mip(velocity, 0, x, 2), make_mip(out_normal, 0, x, 3), make_mip(out_velocity, 0, x, 4));
    else
      hiz_tile_kernel<false><<<grid, 256, 0, st>>>(L, cur, levels, none, none, none, none);
    int rc = check_launch("hiz_tile_kernel");
    if (rc) return rc;
    payload = false;
    cur += levels;
    if (max_levels > 0) break;
  }
  return VKR_OK;
}

static int check_downsample_args(const VkrImage* depth, const VkrImage* normal, const VkrImage* velocity,
                                 const VkrImage* out_normal, const VkrImage* out_velocity) {
  if (!image_ok(depth, VKR_FORMAT_D24_UNORM_S8_UINT) || !image_ok(normal, VKR_FORMAT_R16G16_UNORM) ||
      !image_ok(velocity, VKR_FORMAT_R16G16_SFLOAT) || !image_ok(out_normal, VKR_FORMAT_R16G16_UNORM) ||
      !image_ok(out_velocity, VKR_FORMAT_R16G16_SFLOAT)) {
    set_error("downsample: null image or unexpected format");
    return VKR_ERR_INVALID_ARGUMENT;
  }
  if (depth->mip_count < 2) {  // reference src/downsample_pass.cpp:37-39
    set_error("Can't downsample depth texture with 1 mip level");
    return VKR_ERR_INVALID_ARGUMENT;
  }
  uint32_t w = depth->width / 2 ? depth->width / 2 : 1, h = depth->height / 2 ? depth->height / 2 : 1;
  if (out_normal->width != w || out_normal->height != h || out_velocity->width != w || out_velocity->height != h) {
    set_error("Output textures have different sizes");  // reference src/downsample_pass.cpp:48-50
    return VKR_ERR_SIZE_MISMATCH;
  }
  return VKR_OK;
}

}  // namespace vkr

using namespace vkr;

extern "C" int vkr_downsample_gbuffer(const VkrImage* depth, const VkrImage* normal, const VkrImage* velocity,
                                      const VkrImage* out_normal, const VkrImage* out_velocity, void* stream) {
  int rc = check_downsample_args(depth, normal, velocity, out_normal, out_velocity);
  if (rc) return rc;
  return hiz_run(depth, 0, normal, velocity, out_normal, out_velocity, 1, (cudaStream_t)stream);
}

extern "C" int vkr_depth_mips(const VkrImage* depth, uint32_t src_mip, void* stream) {
  if (!image_ok(depth, VKR_FORMAT_D24_UNORM_S8_UINT) || src_mip >= depth->mip_count) {
    set_error("depth_mips: bad depth image or source mip");
    return VKR_ERR_INVALID_ARGUMENT;
  }
  return hiz_run(depth, (int)src_mip, nullptr, nullptr, nullptr, nullptr, 0, (cudaStream_t)stream);
}

extern "C" int vkr_hiz_build_ex(const VkrImage* depth, const VkrImage* normal, const VkrImage* velocity,
                                const VkrImage* out_normal, const VkrImage* out_velocity, const VkrLaunchExt* ext, void* stream) {
  int rc = check_downsample_args(depth, normal, velocity, out_normal, out_velocity);
  if (rc) return rc;
  if (ext && (ext->row_begin != 0 || ext->row_end != 0)) {
    set_error("hiz_build: row windows are not supported (the pyramid is built whole)");
    return VKR_ERR_UNSUPPORTED;
  }
  ExtInfo x;
  VkrLaunchExt e2;
  if (ext) { e2 = *ext; if (e2.layers > 1 && !e2.params_scratch) e2.params_scratch = (void*)depth->base; }   // no params: scratch unused
  if ((rc = resolve_ext(ext ? &e2 : nullptr, depth->height, 16, &x, "hiz_build"))) return rc;
  if (x.prepare_only()) return VKR_OK;
  return hiz_run(depth, 0, normal, velocity, out_normal, out_velocity, 0, (cudaStream_t)stream, x);
}

extern "C" int vkr_hiz_build(const VkrImage* depth, const VkrImage* normal, const VkrImage* velocity,
                             const VkrImage* out_normal, const VkrImage* out_velocity, void* stream) {
  return vkr_hiz_build_ex(depth, normal, velocity, out_normal, out_velocity, nullptr, stream);
}
