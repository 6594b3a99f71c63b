// passes_ssr_taa.cpp — AdvancedSSR, DeferedShadingPass, TAA task bodies (see passes.hpp).
#include <cuda_runtime.h>

#include <algorithm>

#include "passes.hpp"

using namespace rendergraph;

static void copy_mat(float* dst, const glm::mat4& m) { std::memcpy(dst, m.m, sizeof(m.m)); }

#define HALTON_SEQ_SIZE 128

std::vector<glm::vec4> halton23_seq(uint32_t count) {                                                           // reference src/advanced_ssr.cpp:22-34
  std::vector<float> raw(4 * count);
  vkr_halton23_seq(count, raw.data());
  std::vector<glm::vec4> seq(count);
  for (uint32_t i = 0; i < count; i++) { seq[i].x = raw[4 * i]; seq[i].y = raw[4 * i + 1]; }
  return seq;
}

// ---- AdvancedSSR (reference src/advanced_ssr.cpp) ------------------------------------------------
AdvancedSSR::AdvancedSSR(RenderGraph& graph, uint32_t w, uint32_t h) {
  auto halton_samples = halton23_seq(HALTON_SEQ_SIZE);                                                           // :56-60
  const uint64_t bytes = sizeof(halton_samples[0]) * HALTON_SEQ_SIZE;
  if (cudaMalloc(&halton_buffer, bytes) != cudaSuccess) throw std::runtime_error{"AdvancedSSR: cudaMalloc(halton) failed"};
  if (cudaMemcpy(halton_buffer, halton_samples.data(), bytes, cudaMemcpyHostToDevice) != cudaSuccess) throw std::runtime_error{"AdvancedSSR: halton upload failed"};

  auto mk = [&](uint32_t fmt, uint32_t ww, uint32_t hh) { gpu::ImageInfo i; i.format = fmt; i.width = ww; i.height = hh; return graph.create_image(i); };
  rays = mk(VKR_FORMAT_R16G16B16A16_UNORM, w / 2, h / 2);                                                        // :62-63
  reflections = mk(VKR_FORMAT_R8G8B8A8_UNORM, w / 2, h / 2);                                                     // :68-69
  blurred_reflection = mk(VKR_FORMAT_R8G8B8A8_UNORM, w / 2, h / 2);                                              // :71-73
  blurred_reflection_history = mk(VKR_FORMAT_R8G8B8A8_UNORM, w / 2, h / 2);
  preintegrated_pdf = mk(VKR_FORMAT_R32_SFLOAT, 1024, 1024);                                                     // :88-92
  preintegrated_brdf = mk(VKR_FORMAT_R16G16_SFLOAT, 1024, 1024);

  reflective_indirect = graph.create_buffer(3 * sizeof(uint32_t));                                               // :77-79 VkDispatchIndirectCommand
  glossy_indirect = graph.create_buffer(3 * sizeof(uint32_t));
  const uint64_t tile_bytes = sizeof(int) * std::max<uint64_t>(1, (uint64_t)w * h / 64);                         // :81-83
  reflective_tiles = graph.create_buffer(tile_bytes);
  glossy_tiles = graph.create_buffer(tile_bytes);
  tile_planes = mk(VKR_FORMAT_R32G32B32A32_SFLOAT, std::max(1u, w / 16), std::max(1u, h / 16));                  // :85-86
}

AdvancedSSR::~AdvancedSSR() { cudaFree(halton_buffer); }

void AdvancedSSR::preintegrate_pdf(RenderGraph& graph) {                                                         // :95-114
  struct Input { ImageViewId out_pdf; };
  graph.add_task<Input>("SSR_preintegrate",
    [&](Input& input, RenderGraphBuilder& builder) { input.out_pdf = builder.use_storage_image(preintegrated_pdf, 0, 0); },
    [=](Input& input, RenderResources& resources, gpu::CmdContext& cmd) {
      vkr_check(vkr_pdf_preintegrate(resources.get_image(input.out_pdf), cmd.stream), "SSR_preintegrate");
    });
}

void AdvancedSSR::preintegrate_brdf(RenderGraph& graph) {                                                        // :116-136
  struct Input { ImageViewId out_brdf; };
  const float* halton = halton_buffer;
  graph.add_task<Input>("BRDF_preintegrate",
    [&](Input& input, RenderGraphBuilder& builder) { input.out_brdf = builder.use_storage_image(preintegrated_brdf, 0, 0); },
    [=](Input& input, RenderResources& resources, gpu::CmdContext& cmd) {
      vkr_check(vkr_brdf_preintegrate(halton, resources.get_image(input.out_brdf), cmd.stream), "BRDF_preintegrate");
    });
}

static VkrTraceParams make_trace_params(const AdvancedSSRParams& params, uint32_t counter) {                     // TraceParams, :138-145
  VkrTraceParams c{};
  copy_mat(c.normal_mat, params.normal_mat);
  c.frame_random = counter;
  c.fovy = params.fovy; c.aspect = params.aspect; c.znear = params.znear; c.zfar = params.zfar;
  return c;
}

void AdvancedSSR::run_trace_pass(RenderGraph& graph, const AdvancedSSRParams& params, const Gbuffer& gbuff, ImageResourceId ssr_occlusion) {
  VkrTraceParams config = make_trace_params(params, counter);
  config.max_roughness = settings.max_rougness;                                                                  // push constant, :159-163
  if (settings.update_random) {                                                                                  // :166-171
    counter++;
    counter = counter % settings.max_accumulated_rays;
  }
  struct Input { ImageViewId depth, normal, material, out, occlusion, preintegrated_pdf; };
  auto mips_count = graph.get_descriptor(gbuff.depth).mip_levels;
  const float* halton = halton_buffer;
  graph.add_task<Input>("SSSR_trace",
    [&](Input& input, RenderGraphBuilder& builder) {
      input.depth = builder.sample_image(gbuff.depth, 1, mips_count - 1, 0, 1);                                  // :186
      input.normal = builder.sample_image(gbuff.downsampled_normals);
      input.material = builder.sample_image(gbuff.material);
      input.out = builder.use_storage_image(rays, 0, 0);
      input.occlusion = builder.use_storage_image(ssr_occlusion, 0, 0);
      input.preintegrated_pdf = builder.sample_image(preintegrated_pdf);
    },
    [=](Input& input, RenderResources& resources, gpu::CmdContext& cmd) {
      vkr_check(vkr_sssr_trace(&config, resources.get_image(input.depth), resources.get_image(input.normal), resources.get_image(input.material), halton,
                               resources.get_image(input.preintegrated_pdf), resources.get_image(input.out), resources.get_image(input.occlusion),
                               cmd.stream), "SSSR_trace");
    });
}

void AdvancedSSR::run_filter_pass(RenderGraph& graph, const AdvancedSSRParams& params, const Gbuffer& gbuff) {
  VkrTraceParams config = make_trace_params(params, counter);
  config.render_flags = 0u;                                                                                      // :335-338
  config.render_flags |= settings.normalize_reflections ? VKR_SSR_NORMALIZE_REFLECTIONS : 0;
  config.render_flags |= settings.accumulate_reflections ? VKR_SSR_ACCUMULATE_REFLECTIONS : 0;
  config.render_flags |= settings.bilateral_filter ? VKR_SSR_BILATERAL_FILTER : 0;
  struct Input { ImageViewId depth, normal, albedo, material, rays, reflection; };
  // The reference hard-codes a 10-mip depth view (:342); a 256x256 probe face has 9, so clamp.
  const uint32_t view_mips = std::min<uint32_t>(10u, graph.get_descriptor(gbuff.depth).mip_levels);
  graph.add_task<Input>("SSSR_filter",
    [&](Input& input, RenderGraphBuilder& builder) {
      input.depth = builder.sample_image(gbuff.depth, 0, view_mips, 0, 1);
      input.normal = builder.sample_image(gbuff.normal);
      input.albedo = builder.sample_image(gbuff.albedo);
      input.rays = builder.sample_image(rays);
      input.material = builder.sample_image(gbuff.material);
      input.reflection = builder.use_storage_image(reflections, 0, 0);
    },
    [=](Input& input, RenderResources& resources, gpu::CmdContext& cmd) {
      vkr_check(vkr_sssr_filter(&config, resources.get_image(input.rays), resources.get_image(input.depth), resources.get_image(input.albedo),
                                resources.get_image(input.normal), resources.get_image(input.material), resources.get_image(input.reflection), cmd.stream),
                "SSSR_filter");
    });
}

void AdvancedSSR::run_blur_pass(RenderGraph& graph, const AdvancedSSRParams&, const DrawTAAParams& taa_params, const Gbuffer& gbuff) {
  struct Input { ImageViewId depth, normal, material, reflections, history, velocity, history_depth, result; };
  VkrBlurParams p{};
  p.max_roughness = settings.max_rougness; p.accumulate = settings.accumulate_reflections; p.disable_blur = !settings.use_blur;  // :400
  copy_mat(p.inverse_camera, glm::inverse(taa_params.camera));                                                   // :401
  copy_mat(p.prev_inverse_camera, glm::inverse(taa_params.prev_camera));
  std::memcpy(p.fovy_aspect_znear_zfar, &taa_params.fovy_aspect_znear_zfar, 16);
  const uint32_t view_mips = std::min<uint32_t>(10u, graph.get_descriptor(gbuff.depth).mip_levels);
  graph.add_task<Input>("SSSR_blur",
    [&](Input& input, RenderGraphBuilder& builder) {
      input.depth = builder.sample_image(gbuff.depth, 0, view_mips, 0, 1);
      input.normal = builder.sample_image(gbuff.normal);
      input.reflections = builder.sample_image(reflections);
      input.material = builder.sample_image(gbuff.material);
      input.history = builder.sample_image(blurred_reflection_history);
      input.velocity = builder.sample_image(gbuff.downsampled_velocity_vectors);
      input.history_depth = builder.sample_image(gbuff.prev_depth, 0, view_mips, 0, 1);
      input.result = builder.use_storage_image(blurred_reflection, 0, 0);
    },
    [=](Input& input, RenderResources& resources, gpu::CmdContext& cmd) {
      vkr_check(vkr_sssr_blur(&p, resources.get_image(input.depth), resources.get_image(input.normal), resources.get_image(input.reflections),
                              resources.get_image(input.material), resources.get_image(input.history), resources.get_image(input.velocity),
                              resources.get_image(input.history_depth), resources.get_image(input.result), cmd.stream), "SSSR_blur");
    });
}

// ---- the tile path --------------------------------------------------------------------------------
void AdvancedSSR::clear_indirect_params(RenderGraph& graph) {                                                    // :440-452
  struct Input {};
  auto ri = reflective_indirect, gi = glossy_indirect;
  graph.add_task<Input>("SSSR_Clear",
    [&](Input&, RenderGraphBuilder& builder) {
      builder.transfer_write(reflective_indirect);
      builder.transfer_write(glossy_indirect);
    },
    [=](Input&, RenderResources& resources, gpu::CmdContext& cmd) {
      vkr_check(vkr_sssr_clear_indirect(resources.get_buffer(ri), resources.get_buffer(gi), cmd.stream), "SSSR_Clear");
    });
}

void AdvancedSSR::run_classification_pass(RenderGraph& graph, const AdvancedSSRParams&, const Gbuffer& gbuff) { // :454-495
  struct Input { ImageViewId material_tex; };
  auto extent = graph.get_descriptor(rays).extent2D();
  VkrClassifyParams pc{int(extent.width), int(extent.height), settings.max_rougness, settings.glossy_roughness_value};
  auto ri = reflective_indirect, gi = glossy_indirect, rt = reflective_tiles, gt = glossy_tiles;
  graph.add_task<Input>("SSSR_Classification",
    [&](Input& input, RenderGraphBuilder& builder) {
      input.material_tex = builder.sample_image(gbuff.material);
      builder.use_storage_buffer(reflective_indirect, false);
      builder.use_storage_buffer(glossy_indirect, false);
      builder.use_storage_buffer(reflective_tiles, false);
      builder.use_storage_buffer(glossy_tiles, false);
    },
    [=](Input& input, RenderResources& resources, gpu::CmdContext& cmd) {
      vkr_check(vkr_sssr_classification(&pc, resources.get_image(input.material_tex), (int32_t*)resources.get_buffer(rt), (int32_t*)resources.get_buffer(gt),
                                        resources.get_buffer(ri), resources.get_buffer(gi), cmd.stream), "SSSR_Classification");
    });
}

void AdvancedSSR::run_tile_regression_pass(RenderGraph& graph, const AdvancedSSRParams& params, const Gbuffer& gbuff) {  // :497-538
  struct Input { ImageViewId depth_tex, planes_tex; };
  VkrRegressionParams pc{};
  copy_mat(pc.camera_to_world, glm::transpose(params.normal_mat));                                               // :518
  pc.fovy = params.fovy; pc.aspect = params.aspect; pc.znear = params.znear; pc.zfar = params.zfar;
  graph.add_task<Input>("SSSR_Tile_Regression",
    [&](Input& input, RenderGraphBuilder& builder) {
      input.depth_tex = builder.sample_image(gbuff.depth, 1, 1, 0, 1);                                           // :524
      input.planes_tex = builder.use_storage_image(tile_planes, 0, 0);
    },
    [=](Input& input, RenderResources& resources, gpu::CmdContext& cmd) {
      vkr_check(vkr_sssr_tile_regression(&pc, resources.get_image(input.depth_tex), resources.get_image(input.planes_tex), cmd.stream), "SSSR_Tile_Regression");
    });
}

void AdvancedSSR::run_trace_indirect_pass(RenderGraph& graph, const AdvancedSSRParams& params, const Gbuffer& gbuff) {   // :216-302
  VkrTraceParams config = make_trace_params(params, counter);
  config.max_roughness = settings.max_rougness;                                                                  // :238
  if (settings.update_random) {                                                                                  // :240-243
    counter++;
    counter = counter % settings.max_accumulated_rays;
  }
  struct Input { ImageViewId depth, normal, material, out; };
  auto mips_count = graph.get_descriptor(gbuff.depth).mip_levels;
  const float* halton = halton_buffer;
  auto ri = reflective_indirect, gi = glossy_indirect, rt = reflective_tiles, gt = glossy_tiles;
  const uint32_t max_tiles = (uint32_t)(graph.buffer_size(reflective_tiles) / sizeof(int));
  graph.add_task<Input>("SSSR_trace",
    [&](Input& input, RenderGraphBuilder& builder) {
      input.depth = builder.sample_image(gbuff.depth, 1, mips_count - 1, 0, 1);                                  // :254
      input.normal = builder.sample_image(gbuff.downsampled_normals);
      input.material = builder.sample_image(gbuff.material);
      input.out = builder.use_storage_image(rays, 0, 0);
      builder.use_indirect_buffer(reflective_indirect);
      builder.use_indirect_buffer(glossy_indirect);
      builder.use_storage_buffer(reflective_tiles);
      builder.use_storage_buffer(glossy_tiles);
    },
    [=](Input& input, RenderResources& resources, gpu::CmdContext& cmd) {
      // two dispatch_indirect of one pipeline (:291-299): mirror list, then glossy list
      vkr_check(vkr_sssr_trace_indirect(&config, 0, resources.get_image(input.depth), resources.get_image(input.normal), resources.get_image(input.material),
                                        halton, (const int32_t*)resources.get_buffer(rt), resources.get_buffer(ri), max_tiles, resources.get_image(input.out),
                                        cmd.stream), "SSSR_trace(mirror tiles)");
      vkr_check(vkr_sssr_trace_indirect(&config, 1, resources.get_image(input.depth), resources.get_image(input.normal), resources.get_image(input.material),
                                        halton, (const int32_t*)resources.get_buffer(gt), resources.get_buffer(gi), max_tiles, resources.get_image(input.out),
                                        cmd.stream), "SSSR_trace(glossy tiles)");
    });
}

void AdvancedSSR::run(RenderGraph& graph, const AdvancedSSRParams& params, const DrawTAAParams& taa_params, const Gbuffer& gbuff, ImageResourceId ssr_occlusion) {
  if (settings.run_tile_path) {                                                                                  // :547-550, commented out in the reference
    clear_indirect_params(graph);
    run_classification_pass(graph, params, gbuff);
    run_tile_regression_pass(graph, params, gbuff);
    run_trace_indirect_pass(graph, params, gbuff);
  }
  run_trace_pass(graph, params, gbuff, ssr_occlusion);                                                           // :540-553
  run_filter_pass(graph, params, gbuff);
  run_blur_pass(graph, params, taa_params, gbuff);
}

// ---- DeferedShadingPass (reference src/defered_shading.cpp) ---------------------------------------
DeferedShadingPass::DeferedShadingPass(RenderGraph&) {}

void DeferedShadingPass::update_params(const glm::mat4& camera, const glm::mat4& shadow, float fovy, float aspect, float znear, float zfar) {  // :33-45
  copy_mat(consts.inverse_camera, glm::inverse(camera));
  copy_mat(consts.camera, camera);
  copy_mat(consts.shadow_mvp, shadow);
  consts.fovy = fovy; consts.aspect = aspect; consts.znear = znear; consts.zfar = zfar;
}

void DeferedShadingPass::draw(RenderGraph& graph, const Gbuffer& gbuffer, ImageResourceId, ImageResourceId ssao, ImageResourceId brdf_tex,
                              ImageResourceId reflections, ImageResourceId out_image) {
  struct PassData { ImageViewId albedo, normal, material, depth, rt, ssao, ssr, brdf; };
  VkrShadingParams p = consts;
  p.min_max_roughness[0] = min_max_roughness.x; p.min_max_roughness[1] = min_max_roughness.y;                    // :72
  p.show_ao = only_ao ? 1u : 0u;
  graph.add_task<PassData>("DeferedShading",
    [&](PassData& input, RenderGraphBuilder& builder) {
      input.albedo = builder.sample_image(gbuffer.albedo);
      input.normal = builder.sample_image(gbuffer.normal);
      input.material = builder.sample_image(gbuffer.material);
      input.depth = builder.sample_image(gbuffer.depth);
      input.rt = builder.use_color_attachment(out_image, 0, 0);
      input.ssao = builder.sample_image(ssao);
      input.ssr = builder.sample_image(reflections);
      input.brdf = builder.sample_image(brdf_tex);
    },
    [=](PassData& input, RenderResources& resources, gpu::CmdContext& cmd) {
      vkr_check(vkr_deferred_shading(&p, resources.get_image(input.albedo), resources.get_image(input.normal), resources.get_image(input.material),
                                     resources.get_image(input.depth), resources.get_image(input.ssao), resources.get_image(input.brdf),
                                     resources.get_image(input.ssr), resources.get_image(input.rt), cmd.stream), "DeferedShading");
    });
}

// ---- TAA (reference src/taa.cpp) -------------------------------------------------------------------
TAA::TAA(RenderGraph& graph, uint32_t w, uint32_t h) {
  gpu::ImageInfo info; info.format = VKR_FORMAT_R16G16B16A16_SFLOAT; info.width = w; info.height = h;            // :6
  history = graph.create_image(info);
  target = graph.create_image(info);
}

void TAA::run(RenderGraph& graph, const Gbuffer& gbuffer, ImageResourceId color, const DrawTAAParams& params) {
  struct PassData { ImageViewId history_color, history_depth, current_depth, velocity, color, out; };
  VkrTaaParams consts{};                                                                                          // :30
  copy_mat(consts.inverse_camera, glm::inverse(params.camera));
  copy_mat(consts.prev_inverse_camera, glm::inverse(params.prev_camera));
  std::memcpy(consts.fovy_aspect_znear_zfar, &params.fovy_aspect_znear_zfar, 16);
  graph.add_task<PassData>("TAA",
    [&](PassData& input, RenderGraphBuilder& builder) {
      input.history_color = builder.sample_image(history);
      input.history_depth = builder.sample_image(gbuffer.prev_depth);
      input.current_depth = builder.sample_image(gbuffer.depth);
      input.velocity = builder.sample_image(gbuffer.velocity_vectors);
      input.color = builder.sample_image(color);
      input.out = builder.use_storage_image(target, 0, 0);
    },
    [=](PassData& input, RenderResources& resources, gpu::CmdContext& cmd) {
      vkr_check(vkr_taa_resolve(&consts, resources.get_image(input.history_color), resources.get_image(input.history_depth),
                                resources.get_image(input.current_depth), resources.get_image(input.velocity), resources.get_image(input.color),
                                resources.get_image(input.out), cmd.stream), "TAA");
    });
}
