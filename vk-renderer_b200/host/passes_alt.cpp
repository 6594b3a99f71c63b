// passes_alt.cpp — SSAOPass and add_ssr_pass task bodies (see passes.hpp): the renderer's earlier AO / reflection passes.
#include <cmath>

#include "passes.hpp"

using namespace rendergraph;

static void copy_mat(float* dst, const glm::mat4& m) { std::memcpy(dst, m.m, sizeof(m.m)); }

// ---- SSAO (reference src/ssao.cpp) ----------------------------------------------------------------------------
ImageResourceId create_ssao_texture(RenderGraph& graph, uint32_t width, uint32_t height) {                        // :8-11
  gpu::ImageInfo info; info.format = VKR_FORMAT_R8_UNORM; info.width = width; info.height = height;
  return graph.create_image(info);
}

constexpr uint32_t SAMPLES_COUNT = 16;

namespace {
struct Lcg {   // 64-bit LCG (Knuth's MMIX constants), upper bits: stands in for libc rand() so the set is reproducible
  uint64_t s;
  explicit Lcg(uint32_t seed) : s{0x9E3779B97F4A7C15ull ^ ((uint64_t)seed << 17)} { next(); next(); }
  uint32_t next() { s = s * 6364136223846793005ull + 1442695040888963407ull; return (uint32_t)(s >> 33); }   // 31 bits like RAND_MAX
  float unit() { return 2.f * (float)next() / 2147483647.0f - 1.f; }                                           // :36-38
};
}  // namespace

SSAOPass::SSAOPass(RenderGraph& graph, ImageResourceId target, uint32_t seed) {
  if (graph.get_descriptor(target).format != VKR_FORMAT_R8_UNORM) throw std::runtime_error{"SSAOPass: target must be R8_UNORM"};
  Lcg rng{seed};
  for (uint32_t i = 0; i < SAMPLES_COUNT; i++) {                                                                 // :34-52
    while (true) {
      float x = rng.unit(), y = rng.unit(), z = rng.unit();
      float l2 = x * x + y * y + z * z;
      if (l2 < 1.0f && l2 > 0.0f) {
        float len = std::sqrt(l2);
        sphere_samples.push_back(vec3{x / len, y / len, z / len});
        break;
      }
    }
  }
}

void SSAOPass::draw(RenderGraph& graph, ImageResourceId depth, ImageResourceId target, const SSAOInParams& params) {
  struct PassData { ImageViewId depth, rt; };
  VkrSsaoParams block{};                                                                                          // :66-76
  copy_mat(block.projection, params.projection);
  block.fovy = params.fovy; block.aspect = params.aspect; block.znear = params.znear; block.zfar = params.zfar;
  if (reference_packing) {
    // glm::vec3 samples[16] packed at 12 bytes in the C++ block, read by the shader at 16: 48 floats seen as 12 vec4 slots
    float* flat = &block.samples[0][0];
    for (uint32_t i = 0; i < SAMPLES_COUNT; i++) { flat[3 * i] = sphere_samples[i].x; flat[3 * i + 1] = sphere_samples[i].y; flat[3 * i + 2] = sphere_samples[i].z; }
  } else {
    for (uint32_t i = 0; i < SAMPLES_COUNT; i++) { block.samples[i][0] = sphere_samples[i].x; block.samples[i][1] = sphere_samples[i].y; block.samples[i][2] = sphere_samples[i].z; }
  }
  last_block = block;
  graph.add_task<PassData>("SSAO",
    [&](PassData& input, RenderGraphBuilder& builder) {
      input.depth = builder.sample_image(depth, 0, 1, 0, 1);                                                      // :62
      input.rt = builder.use_color_attachment(target, 0, 0);
    },
    [=](PassData& input, RenderResources& resources, gpu::CmdContext& cmd) {
      vkr_check(vkr_ssao(&block, resources.get_image(input.depth), resources.get_image(input.rt), cmd.stream), "SSAO");
    });
}

// ---- SSR (reference src/ssr.cpp) -------------------------------------------------------------------------------
ImageResourceId create_ssr_tex(RenderGraph& graph, uint32_t w, uint32_t h) {                                      // :5-8
  gpu::ImageInfo info; info.format = VKR_FORMAT_R8G8B8A8_UNORM; info.width = w; info.height = h;
  return graph.create_image(info);
}

void add_ssr_pass(RenderGraph& graph, ImageResourceId depth, ImageResourceId normal, ImageResourceId color, ImageResourceId material,
                  ImageResourceId out, const SSRParams& params) {
  struct Input { ImageViewId depth, normal, color, material, rt; };
  VkrSsrParams block{};                                                                                           // :52-53
  copy_mat(block.normal_mat, params.normal_mat);
  block.fovy = params.fovy; block.aspect = params.aspect; block.znear = params.znear; block.zfar = params.zfar;
  graph.add_task<Input>("SSR",
    [&](Input& input, RenderGraphBuilder& builder) {
      input.depth = builder.sample_image(depth);
      input.normal = builder.sample_image(normal);
      input.color = builder.sample_image(color);
      input.material = builder.sample_image(material);
      input.rt = builder.use_color_attachment(out, 0, 0);
    },
    [=](Input& input, RenderResources& resources, gpu::CmdContext& cmd) {
      vkr_check(vkr_ssr_legacy(&block, resources.get_image(input.depth), resources.get_image(input.normal), resources.get_image(input.color),
                               resources.get_image(input.material), resources.get_image(input.rt), cmd.stream), "SSR");
    });
}

// ---- ScreenSpaceTrace (reference src/screen_trace.cpp) ----------------------------------------------------------
ScreenSpaceTrace::ScreenSpaceTrace(RenderGraph& graph, uint32_t width, uint32_t height) {                         // :3-21
  gpu::ImageInfo info; info.format = VKR_FORMAT_R16G16B16A16_SFLOAT; info.width = width; info.height = height;
  raw = graph.create_image(info);
  filtered = graph.create_image(info);
  accumulated = graph.create_image(info);
}

void ScreenSpaceTrace::add_main_pass(RenderGraph& graph, const ScreenTraceParams& params, ImageResourceId depth, ImageResourceId normal,
                                     ImageResourceId color, ImageResourceId material) {
  struct PassData { ImageViewId out, depth, norm, color, material; };
  const float angle_offsets[]{60.f, 300.f, 180.f, 240.f, 120.f, 0.f, 300.f, 60.f, 180.f, 120.f, 240.f, 0.f};       // :47-54
  float base_angle = angle_offsets[frame_count % (sizeof(angle_offsets) / sizeof(float))] / 360.f;
  base_angle += next_angle_random - 0.5f;
  frame_count += 1;
  VkrScreenTraceParams block{};
  copy_mat(block.normal_mat, params.normal_mat);
  block.random_offset = next_random_offset; block.angle_offset = base_angle;
  block.fovy = params.fovy; block.aspect = params.aspect; block.znear = params.znear; block.zfar = params.zfar;
  last_block = block;
  graph.add_task<PassData>("ScreenTrace",
    [&](PassData& input, RenderGraphBuilder& builder) {
      input.depth = builder.sample_image(depth, 0, 1, 0, 1);
      input.norm = builder.sample_image(normal);
      input.color = builder.sample_image(color);
      input.material = builder.sample_image(material);
      input.out = builder.use_storage_image(raw, 0, 0);
    },
    [=](PassData& input, RenderResources& resources, gpu::CmdContext& cmd) {
      vkr_check(vkr_screen_trace_main(&block, resources.get_image(input.depth), resources.get_image(input.norm), resources.get_image(input.color),
                                      resources.get_image(input.material), resources.get_image(input.out), cmd.stream), "ScreenTrace");
    });
}

void ScreenSpaceTrace::add_filter_pass(RenderGraph& graph, const ScreenTraceParams& params, ImageResourceId depth) {
  struct PassData { ImageViewId depth, raw, filtered; };
  const float znear = params.znear, zfar = params.zfar;                                                            // :117
  graph.add_task<PassData>("ScreenTraceFilter",
    [&](PassData& input, RenderGraphBuilder& builder) {
      input.depth = builder.sample_image(depth, 0, 1, 0, 1);
      input.raw = builder.sample_image(raw);
      input.filtered = builder.use_storage_image(filtered, 0, 0);
    },
    [=](PassData& input, RenderResources& resources, gpu::CmdContext& cmd) {
      vkr_check(vkr_screen_trace_filter(znear, zfar, resources.get_image(input.raw), resources.get_image(input.depth), resources.get_image(input.filtered),
                                        cmd.stream), "ScreenTraceFilter");
    });
}

void ScreenSpaceTrace::add_accumulate_pass(RenderGraph& graph, const ScreenTraceParams& params, ImageResourceId depth, ImageResourceId prev_depth) {
  struct PassData { ImageViewId depth, prev_depth, filtered, accum; };
  struct Pc { float v[4]; } pc{{params.fovy, params.aspect, params.znear, params.zfar}};                           // :160
  graph.add_task<PassData>("ScreenTraceAccumulate",
    [&](PassData& input, RenderGraphBuilder& builder) {
      input.depth = builder.sample_image(depth, 0, 1, 0, 1);
      input.prev_depth = builder.sample_image(prev_depth, 0, 1, 0, 1);
      input.filtered = builder.sample_image(filtered);
      input.accum = builder.use_storage_image(accumulated, 0, 0);
    },
    [=](PassData& input, RenderResources& resources, gpu::CmdContext& cmd) {
      vkr_check(vkr_screen_trace_accumulate(pc.v, resources.get_image(input.depth), resources.get_image(input.prev_depth), resources.get_image(input.filtered),
                                            resources.get_image(input.accum), cmd.stream), "ScreenTraceAccumulate");
    });
}
