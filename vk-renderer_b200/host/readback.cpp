// readback.cpp — see readback.hpp.
#include "readback.hpp"

#include <cuda_runtime.h>

#include <cstdio>
#include <cstring>
#include <fstream>
#include <vector>

using namespace rendergraph;

void blit_image(RenderGraph& graph, ImageResourceId src, ImageResourceId dst) {                                   // util_passes.cpp:152-180
  struct Input {};
  graph.add_task<Input>("CopyImage",
    [&](Input&, RenderGraphBuilder& builder) {
      builder.transfer_read(src, 0, 1, 0, 1);
      builder.transfer_write(dst, 0, 1, 0, 1);
    },
    [=](Input&, RenderResources& resources, gpu::CmdContext& cmd) {
      vkr_check(vkr_blit_image(resources.get_image(src), resources.get_image(dst), cmd.stream), "CopyImage");
    });
}

void ReadBackSystem::release(Request& r) {
  if (r.device) cudaFree(r.device);
  if (r.host) cudaFreeHost(r.host);
  if (r.done) cudaEventDestroy((cudaEvent_t)r.done);
  r.device = r.host = r.done = nullptr;
}

ReadBackSystem::~ReadBackSystem() { clear(); }

void ReadBackSystem::clear() {
  for (auto& kv : requests) {
    if (kv.second.done) cudaEventSynchronize((cudaEvent_t)kv.second.done);
    release(kv.second);
  }
  requests.clear();
  processed_requests.clear();
}

ReadBackID ReadBackSystem::read_image(RenderGraph& graph, ImageResourceId image) { return read_image(graph, image, 0, 0, VKR_READBACK_RAW); }   // :67-69

ReadBackID ReadBackSystem::read_image(RenderGraph& graph, ImageResourceId image, uint32_t mip, uint32_t layer, uint32_t mode) {            // :71-110
  struct TaskData {};
  const auto desc = graph.get_descriptor(image);
  if (mip >= desc.mip_levels || layer >= desc.array_layers) throw std::runtime_error{"Image subresource out of range"};
  const uint32_t image_width = std::max(1u, desc.width / (1u << mip));
  const uint32_t image_height = std::max(1u, desc.height / (1u << mip));
  const uint32_t texel_size = vkr_format_texel_size(desc.format);
  if (!texel_size) throw std::runtime_error{"Unsupported format"};                                                                         // :60-62
  const size_t bytes = (size_t)image_height * image_width * texel_size;

  Request rq{graph.get_frames_count() + 1, image_width, image_height, desc.format, texel_size};
  cudaEvent_t ev;
  if (cudaMalloc(&rq.device, bytes) != cudaSuccess || cudaMallocHost(&rq.host, bytes) != cudaSuccess ||
      cudaEventCreateWithFlags(&ev, cudaEventDisableTiming) != cudaSuccess) {
    release(rq);
    throw std::runtime_error{"ReadBackSystem: buffer allocation failed"};
  }
  rq.done = ev;
  const ReadBackID id = next_request_id++;
  requests[id] = rq;
  const uint64_t layer_off = graph.layer_stride(image) * layer;
  void* dev = rq.device; void* host = rq.host;

  graph.add_task<TaskData>("ImageRead",
    [&](TaskData&, RenderGraphBuilder& builder) { builder.transfer_read(image, mip, 1, layer, 1); },
    [=](TaskData&, RenderResources& resources, gpu::CmdContext& ctx) {
      VkrImage im = *resources.get_image(image);
      im.base = (unsigned char*)im.base + layer_off;
      vkr_check(vkr_readback_pack(&im, mip, mode, dev, ctx.stream), "ImageRead");
      if (cudaMemcpyAsync(host, dev, bytes, cudaMemcpyDeviceToHost, (cudaStream_t)ctx.stream) != cudaSuccess ||
          cudaEventRecord(ev, (cudaStream_t)ctx.stream) != cudaSuccess)
        throw std::runtime_error{"ImageRead: device-to-host copy failed"};
    });
  graph.mark_last_task_not_capturable();   // records an event the host waits on in after_submit(): the frame is issued eagerly
  return id;
}

void ReadBackSystem::after_submit(RenderGraph&) {                                                                                          // :112-146
  auto it = requests.begin();
  while (it != requests.end()) {
    if (it->second.wait_frames > 0) {
      it->second.wait_frames--;
      it++;
      continue;
    }
    auto id = it->first;
    auto& src_request = it->second;
    if (!src_request.host) throw std::runtime_error{"Error, readback got unmapped cpu pointer"};
    if (cudaEventSynchronize((cudaEvent_t)src_request.done) != cudaSuccess) throw std::runtime_error{"ReadBackSystem: copy did not complete"};

    ReadBackData dst_request{};
    dst_request.width = src_request.width;
    dst_request.height = src_request.height;
    dst_request.texel_size = src_request.texel_size;
    dst_request.texel_fmt = src_request.texel_fmt;
    const size_t byte_count = (size_t)dst_request.width * dst_request.height * dst_request.texel_size;
    dst_request.bytes.reset(new uint8_t[byte_count]);
    std::memcpy(dst_request.bytes.get(), src_request.host, byte_count);
    processed_requests[id] = std::move(dst_request);
    release(src_request);
    it = requests.erase(it);
  }
}

ReadBackData ReadBackSystem::get_data(ReadBackID id) {                                                                                     // :148-160
  ReadBackData readback{};
  auto& src = processed_requests.at(id);
  std::swap(readback.bytes, src.bytes);
  std::swap(readback.width, src.width);
  std::swap(readback.height, src.height);
  std::swap(readback.texel_fmt, src.texel_fmt);
  std::swap(readback.texel_size, src.texel_size);
  processed_requests.erase(id);
  return readback;
}

// ---- serialisers ----------------------------------------------------------------------------------------------------------
bool get_depth_cb(ReadBackData&& image, const char* save_path) {                                                                          // main.cpp:118-150
  const uint32_t* word_ptr = reinterpret_cast<const uint32_t*>(image.bytes.get());
  std::ofstream file{save_path, std::ios::trunc};
  if (!file) return false;
  file << "y, ";
  for (uint32_t x = 0; x < image.width; x++) {
    file << x;
    if (x != image.width - 1) file << ",";
  }
  file << "\n";
  for (uint32_t y = 0; y < image.height; y++) {
    file << y << ",";
    file << std::hex;
    for (uint32_t x = 0; x < image.width; x++) {
      auto depth = 0xffffff & word_ptr[y * image.width + x];
      file << "0x" << depth;
      if (x != image.width - 1) file << ",";
    }
    file << std::dec;
    file << "\n";
  }
  file.close();
  return (bool)file;
}

bool get_depth_png_cb(ReadBackData&& image, const char* save_path) {                                                                      // main.cpp:152-161
  uint32_t* word_ptr = reinterpret_cast<uint32_t*>(image.bytes.get());
  for (size_t i = 0; i < (size_t)image.width * image.height; i++) word_ptr[i] &= 0xffffff;   // no-op after VKR_READBACK_DEPTH24
  return write_png_rgba8(save_path, image.width, image.height, image.bytes.get());
}

bool get_rgba_cb(ReadBackData&& image, const char* save_path) {                                                                           // main.cpp:163-176
  for (size_t i = 0; i < (size_t)image.width * image.height; i++) image.bytes[4 * i + 3] = 255;   // no-op after VKR_READBACK_OPAQUE_ALPHA
  return write_png_rgba8(save_path, image.width, image.height, image.bytes.get());
}

namespace {
struct Crc {
  uint32_t table[256];
  Crc() {
    for (uint32_t n = 0; n < 256; n++) {
      uint32_t c = n;
      for (int k = 0; k < 8; k++) c = (c & 1) ? 0xEDB88320u ^ (c >> 1) : c >> 1;
      table[n] = c;
    }
  }
  uint32_t update(uint32_t crc, const uint8_t* p, size_t n) const {
    for (size_t i = 0; i < n; i++) crc = table[(crc ^ p[i]) & 255] ^ (crc >> 8);
    return crc;
  }
};
void put32(std::vector<uint8_t>& v, uint32_t x) { v.push_back(x >> 24); v.push_back(x >> 16); v.push_back(x >> 8); v.push_back(x); }
bool chunk(FILE* f, const Crc& crc, const char* type, const uint8_t* data, size_t n) {
  uint8_t hdr[8] = {(uint8_t)(n >> 24), (uint8_t)(n >> 16), (uint8_t)(n >> 8), (uint8_t)n, (uint8_t)type[0], (uint8_t)type[1], (uint8_t)type[2], (uint8_t)type[3]};
  uint32_t c = crc.update(0xFFFFFFFFu, hdr + 4, 4);
  if (n) c = crc.update(c, data, n);
  c ^= 0xFFFFFFFFu;
  const uint8_t tail[4] = {(uint8_t)(c >> 24), (uint8_t)(c >> 16), (uint8_t)(c >> 8), (uint8_t)c};
  return fwrite(hdr, 1, 8, f) == 8 && (n == 0 || fwrite(data, 1, n, f) == n) && fwrite(tail, 1, 4, f) == 4;
}
}  // namespace

bool write_png_rgba8(const char* path, uint32_t width, uint32_t height, const uint8_t* rgba) {
  if (!path || !rgba || !width || !height) return false;
  static const Crc crc;
  // raw scanlines: filter byte 0 + 4 * width bytes
  const size_t stride = (size_t)width * 4 + 1, raw_size = stride * height;
  std::vector<uint8_t> z;
  z.reserve(raw_size + raw_size / 65535 * 5 + 16);
  z.push_back(0x78); z.push_back(0x01);                                   // zlib header, no preset dictionary
  uint32_t s1 = 1, s2 = 0;                                                // Adler-32 of the raw stream
  std::vector<uint8_t> raw(raw_size);
  for (uint32_t y = 0; y < height; y++) {
    raw[y * stride] = 0;
    std::memcpy(&raw[y * stride + 1], rgba + (size_t)y * width * 4, (size_t)width * 4);
  }
  for (size_t off = 0; off < raw_size;) {
    const size_t n = std::min<size_t>(65535, raw_size - off);
    z.push_back(off + n == raw_size ? 1 : 0);                             // BFINAL, BTYPE = 00 (stored)
    z.push_back(n & 255); z.push_back(n >> 8); z.push_back(~n & 255); z.push_back((~n >> 8) & 255);
    z.insert(z.end(), raw.begin() + off, raw.begin() + off + n);
    off += n;
  }
  for (size_t i = 0; i < raw_size;) {                                      // Adler-32 in blocks that cannot overflow
    const size_t n = std::min<size_t>(5552, raw_size - i);
    for (size_t k = 0; k < n; k++) { s1 += raw[i + k]; s2 += s1; }
    s1 %= 65521; s2 %= 65521;
    i += n;
  }
  put32(z, (s2 << 16) | s1);
  FILE* f = fopen(path, "wb");
  if (!f) return false;
  static const uint8_t sig[8] = {0x89, 'P', 'N', 'G', 0x0D, 0x0A, 0x1A, 0x0A};
  std::vector<uint8_t> ihdr;
  put32(ihdr, width); put32(ihdr, height);
  ihdr.push_back(8); ihdr.push_back(6); ihdr.push_back(0); ihdr.push_back(0); ihdr.push_back(0);   // 8 bit, RGBA, deflate, adaptive, no interlace
  bool ok = fwrite(sig, 1, 8, f) == 8 && chunk(f, crc, "IHDR", ihdr.data(), ihdr.size()) && chunk(f, crc, "IDAT", z.data(), z.size()) &&
            chunk(f, crc, "IEND", nullptr, 0);
  ok = (fclose(f) == 0) && ok;
  return ok;
}
