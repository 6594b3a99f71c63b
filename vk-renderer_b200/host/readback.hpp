// readback.hpp — the capture path of the reference's frame loop, headless:
//   blit_image                      reference src/util_passes.hpp / util_passes.cpp:152-180 (task "CopyImage")
//   ReadBackSystem, ReadBackData    reference src/image_readback.hpp:10-49, image_readback.cpp:67-160 (task "ImageRead")
//   get_depth_cb / get_depth_png_cb / get_rgba_cb   reference src/main.cpp:118-176 (CSV of 24-bit depth codes, PNG of the depth
//                                   words, PNG of an RGBA8 frame with alpha forced to 255)
// Same classes and call sequence (read_image -> submit -> after_submit -> is_data_available -> get_data -> callback).  The
// mask / alpha fix-ups of the callbacks run on the device while the level is packed (vkr_readback_pack), so the host-side
// callbacks only serialise; stb_image_write is replaced by a small PNG encoder (stored deflate blocks).
#pragma once
#include <memory>
#include <string>
#include <unordered_map>

#include "rendergraph.hpp"

void blit_image(rendergraph::RenderGraph& graph, rendergraph::ImageResourceId src, rendergraph::ImageResourceId dst);

struct ReadBackData {
  uint32_t width = 0;
  uint32_t height = 0;
  uint32_t texel_fmt = 0;   // VkrFormat
  uint32_t texel_size = 0;
  std::unique_ptr<uint8_t[]> bytes{nullptr};
};

using ReadBackID = uint64_t;
const ReadBackID INVALID_READBACK = ~0ull;

struct ReadBackSystem {
  ~ReadBackSystem();
  ReadBackID read_image(rendergraph::RenderGraph& graph, rendergraph::ImageResourceId image);
  // `mode` (VkrReadbackMode) is not in the reference: the fix-up its callback would apply, done while packing
  ReadBackID read_image(rendergraph::RenderGraph& graph, rendergraph::ImageResourceId image, uint32_t mip, uint32_t layer, uint32_t mode = VKR_READBACK_RAW);
  void after_submit(rendergraph::RenderGraph& graph);
  bool is_data_available(ReadBackID id) const { return processed_requests.count(id); }
  ReadBackData get_data(ReadBackID id);
  void clear();

 private:
  struct Request {
    uint32_t wait_frames;
    uint32_t width, height;
    uint32_t texel_fmt, texel_size;
    void* device = nullptr;   // tight device buffer
    void* host = nullptr;     // pinned
    void* done = nullptr;     // cudaEvent_t recorded behind the device-to-host copy
  };
  ReadBackID next_request_id = 0;
  std::unordered_map<ReadBackID, Request> requests;
  std::unordered_map<ReadBackID, ReadBackData> processed_requests;
  static void release(Request& r);
};

// reference src/main.cpp:118-176.  Each returns false when the file cannot be written.
bool get_depth_cb(ReadBackData&& image, const char* save_path = "captures/gbuffer_depth.csv");
bool get_depth_png_cb(ReadBackData&& image, const char* save_path = "captures/gbuffer_depth.png");
bool get_rgba_cb(ReadBackData&& image, const char* save_path);

// 8-bit RGBA PNG, rows tightly packed (what stbi_write_png(path, w, h, 4, data, 0) writes, uncompressed).
bool write_png_rgba8(const char* path, uint32_t width, uint32_t height, const uint8_t* rgba);
