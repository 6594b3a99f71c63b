// ORACLE — TEST INFRASTRUCTURE ONLY (see oracle.h header).  PARITY UNPINNED.
//
// The capture path behind the reference's "Final frame" / depth buttons:
//   blit_image ("CopyImage")       reference src/util_passes.cpp:152-180 — vkCmdBlitImage, VK_FILTER_LINEAR, whole image to whole image
//   ReadBackSystem::read_image     reference src/image_readback.cpp:67-110 — vkCmdCopyImageToBuffer, tightly packed rows
//   get_depth_png_cb / get_rgba_cb reference src/main.cpp:152-176 — depth words masked to 24 bits, alpha forced to 255
// Vulkan defines the blit, not the reference: destination texel (i, j) samples the source at unnormalised coordinate
// (i + 0.5) * src_w / dst_w with linear filtering and clamp to edge, which is texture(src, (p + 0.5) / dst_size) in the terms of
// App. A.3; the filtered value is converted to the destination format (sRGB-encoded for an _SRGB target).
#include <cstring>
#include "oracle.h"
#include "shader_include.hpp"

using namespace orc;

extern "C" int orc_blit_image(const OrcImage* src_, const OrcImage* dst_) {
  const Image& dst = *(const Image*)dst_;
  View src(*(const Image*)src_, 0, 1);
  const int W = (int)dst.width, H = (int)dst.height;
#pragma omp parallel for schedule(dynamic, 8)
  for (int py = 0; py < H; py++)
    for (int px = 0; px < W; px++) {
      vec2 uv = (vec2((float)px, (float)py) + vec2(0.5f, 0.5f)) / vec2((float)W, (float)H);
      imageStore(dst, 0, ivec2(px, py), texture(src, uv));
    }
  return 0;
}

// mode 0 = bytes as they are, 1 = 32-bit words & 0xFFFFFF (main.cpp:156), 2 = byte 3 of every texel <- 255 (main.cpp:166)
extern "C" int orc_readback_pack(const OrcImage* image_, uint32_t mip, uint32_t mode, void* out_tight) {
  const Image& im = *(const Image*)image_;
  if (mip >= im.mip_count) return -1;
  const int w = mip_dim(im.width, (int)mip), h = mip_dim(im.height, (int)mip);
  const uint32_t ts = texel_size(im.format);
  if (mode != 0 && ts != 4) return -1;
  for (int y = 0; y < h; y++) {
    uint8_t* row = (uint8_t*)out_tight + (size_t)y * w * ts;
    std::memcpy(row, texel_ptr(im, 0, y, (int)mip), (size_t)w * ts);
    if (mode == 1) for (int x = 0; x < w; x++) row[4 * x + 3] = 0;
    if (mode == 2) for (int x = 0; x < w; x++) row[4 * x + 3] = 255;
  }
  return 0;
}
