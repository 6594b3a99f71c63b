// readback.cu — the device half of the capture path:
//   "CopyImage"  blit_image, reference src/util_passes.cpp:152-180: vkCmdBlitImage with VK_FILTER_LINEAR, whole image to whole
//                image (the TAA target RGBA16F -> the RGBA8_SRGB readback image, src/main.cpp:220-223,:394)
//   "ImageRead"  ReadBackSystem::read_image, reference src/image_readback.cpp:67-110: vkCmdCopyImageToBuffer of one mip into a
//                tightly packed buffer; the fix-ups the reference's callbacks do on the host afterwards (depth words masked to
//                24 bits, alpha forced to 255: src/main.cpp:152-176) ride along in the same pass over the bytes.
#include "common.cuh"

namespace vkr {

__device__ __forceinline__ float4 rb_unpack_rgba16f(uint2 k) {
  float2 a = unpack_rg16f(k.x), b = unpack_rg16f(k.y);
  return make_float4(a.x, a.y, b.x, b.y);
}
__device__ __forceinline__ float4 rb_unpack_rgba8(uint32_t k) {
  return make_float4(unorm8_to_float(k & 255u), unorm8_to_float((k >> 8) & 255u), unorm8_to_float((k >> 16) & 255u), unorm8_to_float(k >> 24));
}

enum { SRC_RGBA16F = 0, SRC_SRGB8 = 1, SRC_UNORM8 = 2 };

template <int SRC>
__device__ __forceinline__ float4 rb_fetch(const Mip& m, int x, int y, const float* __restrict__ srgb) {
  if (SRC == SRC_RGBA16F) return rb_unpack_rgba16f(ld64(m, x, y));
  if (SRC == SRC_SRGB8) return unpack_srgb8(ld32(m, x, y), srgb);
  return rb_unpack_rgba8(ld32(m, x, y));
}

// Destination texel p samples the source at (p + 0.5) / dst_size with the sampler of App. A.3 (unfused, like every texel-selecting
// coordinate in this library), then is converted to the destination format.
template <int SRC, bool DST_SRGB>
__global__ void __launch_bounds__(256) blit_kernel(Mip src, Mip dst, Dim dd, const float* __restrict__ srgb) {
  const int px = blockIdx.x * 32 + (threadIdx.x & 31), py = blockIdx.y * 8 + (threadIdx.x >> 5);
  if (px >= dst.w || py >= dst.h) return;
  const Tap t = make_tap(pixel_uv(px, dd.w, dd.rw), pixel_uv(py, dd.h, dd.rh), src.w, src.h);
  const float4 a = rb_fetch<SRC>(src, t.x0, t.y0, srgb), b = rb_fetch<SRC>(src, t.x1, t.y0, srgb);
  const float4 c = rb_fetch<SRC>(src, t.x0, t.y1, srgb), d = rb_fetch<SRC>(src, t.x1, t.y1, srgb);
  const float r = bilerp(a.x, b.x, c.x, d.x, t.wx, t.wy), g = bilerp(a.y, b.y, c.y, d.y, t.wx, t.wy);
  const float bl = bilerp(a.z, b.z, c.z, d.z, t.wx, t.wy), al = bilerp(a.w, b.w, c.w, d.w, t.wx, t.wy);
  const uint32_t o = DST_SRGB ? pack_rgba8unorm(srgb_encode(r), srgb_encode(g), srgb_encode(bl), al) : pack_rgba8unorm(r, g, bl, al);
  *((uint32_t*)(dst.p + (size_t)py * dst.pitch) + px) = o;
}

// Rows of `words_per_row` 32-bit words from a pitched level into a tight buffer; and_mask / or_mask applied to every word.
// VEC: four words per thread through 128-bit accesses (row length a multiple of 16 bytes; pitch and bases are by layout).
template <bool VEC>
__global__ void __launch_bounds__(256) pack_rows_kernel(const unsigned char* __restrict__ src, unsigned pitch, int words_per_row, int rows,
                                                        uint32_t* __restrict__ out, uint32_t and_mask, uint32_t or_mask) {
  const int y = blockIdx.y;
  if (y >= rows) return;
  const int per = VEC ? words_per_row / 4 : words_per_row;
  for (int i = blockIdx.x * 256 + threadIdx.x; i < per; i += gridDim.x * 256) {
    if (VEC) {
      uint4 v = __ldg((const uint4*)(src + (size_t)y * pitch) + i);
      v.x = (v.x & and_mask) | or_mask; v.y = (v.y & and_mask) | or_mask; v.z = (v.z & and_mask) | or_mask; v.w = (v.w & and_mask) | or_mask;
      ((uint4*)(out + (size_t)y * words_per_row))[i] = v;
    } else {
      out[(size_t)y * words_per_row + i] = (__ldg((const uint32_t*)(src + (size_t)y * pitch) + i) & and_mask) | or_mask;
    }
  }
}

}  // namespace vkr
using namespace vkr;

extern "C" int vkr_blit_image(const VkrImage* src, const VkrImage* dst, void* stream) {
  if (!src || !dst || !src->base || !dst->base || !src->width || !src->height || !dst->width || !dst->height) {
    set_error("blit_image: null argument or empty image");
    return VKR_ERR_INVALID_ARGUMENT;
  }
  int s = -1;
  if (src->format == VKR_FORMAT_R16G16B16A16_SFLOAT) s = SRC_RGBA16F;
  else if (src->format == VKR_FORMAT_R8G8B8A8_SRGB) s = SRC_SRGB8;
  else if (src->format == VKR_FORMAT_R8G8B8A8_UNORM) s = SRC_UNORM8;
  const bool dsrgb = dst->format == VKR_FORMAT_R8G8B8A8_SRGB;
  if (s < 0 || (!dsrgb && dst->format != VKR_FORMAT_R8G8B8A8_UNORM)) {
    set_error("blit_image: source must be RGBA16F / RGBA8 (sRGB or UNORM), destination RGBA8 (sRGB or UNORM)");
    return VKR_ERR_UNSUPPORTED;
  }
  const float* srgb = srgb_decode_table_device();
  if (!srgb) { set_error("blit_image: sRGB table upload failed"); return VKR_ERR_CUDA; }
  const Mip sm = make_mip(src, 0), dm = make_mip(dst, 0);
  const Dim dd = make_dim(dst->width, dst->height);
  dim3 grid((dst->width + 31) / 32, (dst->height + 7) / 8);
  cudaStream_t st = (cudaStream_t)stream;
#define BLIT(S)                                                               \
  if (dsrgb) blit_kernel<S, true><<<grid, 256, 0, st>>>(sm, dm, dd, srgb);   \
  else blit_kernel<S, false><<<grid, 256, 0, st>>>(sm, dm, dd, srgb);
  if (s == SRC_RGBA16F) { BLIT(SRC_RGBA16F) } else if (s == SRC_SRGB8) { BLIT(SRC_SRGB8) } else { BLIT(SRC_UNORM8) }
#undef BLIT
  return check_launch("blit_kernel");
}

extern "C" int vkr_readback_pack(const VkrImage* image, uint32_t mip, uint32_t mode, void* out_tight, void* stream) {
  if (!image || !image->base || !out_tight || mip >= image->mip_count || mode > VKR_READBACK_OPAQUE_ALPHA) {
    set_error("readback_pack: null argument, mip out of range or unknown mode");
    return VKR_ERR_INVALID_ARGUMENT;
  }
  const uint32_t ts = vkr_format_texel_size(image->format);
  const Mip m = make_mip(image, mip);
  const size_t row_bytes = (size_t)m.w * ts;
  if (!ts || (mode != VKR_READBACK_RAW && ts != 4) || (row_bytes & 3)) {
    set_error("readback_pack: the fix-up modes need 4-byte texels; rows must be a whole number of 32-bit words");
    return VKR_ERR_UNSUPPORTED;
  }
  const uint32_t and_mask = mode == VKR_READBACK_DEPTH24 ? 0x00FFFFFFu : 0xFFFFFFFFu, or_mask = mode == VKR_READBACK_OPAQUE_ALPHA ? 0xFF000000u : 0u;
  const int words = (int)(row_bytes / 4);
  const bool vec = (row_bytes % 16) == 0 && ((uintptr_t)out_tight % 16) == 0;
  const int per = vec ? words / 4 : words;
  int gx = (per + 255) / 256;
  if (gx > 64) gx = 64;
  if (gx < 1) gx = 1;
  dim3 grid((unsigned)gx, (unsigned)m.h);
  cudaStream_t st = (cudaStream_t)stream;
  if (vec) pack_rows_kernel<true><<<grid, 256, 0, st>>>(m.p, m.pitch, words, m.h, (uint32_t*)out_tight, and_mask, or_mask);
  else pack_rows_kernel<false><<<grid, 256, 0, st>>>(m.p, m.pitch, words, m.h, (uint32_t*)out_tight, and_mask, or_mask);
  return check_launch("pack_rows_kernel");
}
