// passes.cpp — task bodies of the pass classes; see passes.hpp for the reference surfaces.
#include "passes.hpp"

#include <cuda_runtime.h>

#include <algorithm>

#include "../../include/vkr_synth.h"

using namespace rendergraph;

namespace glmlite {
mat4 inverse(const mat4& a) { mat4 r; vkr_mat4_inverse(a.m, r.m); return r; }
mat4 transpose(const mat4& a) { mat4 r; vkr_mat4_transpose(a.m, r.m); return r; }
}  // namespace glmlite

static void copy_mat(float* dst, const glm::mat4& m) { std::memcpy(dst, m.m, sizeof(m.m)); }

// ---- Gbuffer (reference src/scene_renderer.cpp:8-44) ------------------------------------------
Gbuffer::Gbuffer(RenderGraph& graph, uint32_t width, uint32_t height) : w{width}, h{height} {
  const uint32_t mips = vkr_full_mip_count(width, height);   // floor(log2(max(w,h))) + 1, :13
  auto mk = [&](uint32_t fmt, uint32_t ww, uint32_t hh, uint32_t m) { gpu::ImageInfo i; i.format = fmt; i.width = ww; i.height = hh; i.mip_levels = m; return graph.create_image(i); };
  albedo = mk(VKR_FORMAT_R8G8B8A8_SRGB, width, height, 1);
  normal = mk(VKR_FORMAT_R16G16_UNORM, width, height, 1);
  velocity_vectors = mk(VKR_FORMAT_R16G16_SFLOAT, width, height, 1);
  material = mk(VKR_FORMAT_R8G8B8A8_SRGB, width, height, 1);
  depth = mk(VKR_FORMAT_D24_UNORM_S8_UINT, width, height, mips);
  prev_depth = mk(VKR_FORMAT_D24_UNORM_S8_UINT, width, height, mips);
  downsampled_normals = mk(VKR_FORMAT_R16G16_UNORM, width / 2, height / 2, 1);
  downsampled_velocity_vectors = mk(VKR_FORMAT_R16G16_SFLOAT, width / 2, height / 2, 1);
}

// ---- DownsamplePass (reference src/downsample_pass.cpp) -----------------------------------------
void DownsamplePass::run_downsample_gbuff(RenderGraph& graph, ImageResourceId src_normals, ImageResourceId src_velocity, ImageResourceId depth,
                                          ImageResourceId out_normal, ImageResourceId out_velocity) {
  auto depth_desc = graph.get_descriptor(depth);
  auto norm_desc = graph.get_descriptor(out_normal);
  auto vel_desc = graph.get_descriptor(out_velocity);
  if (depth_desc.mip_levels < 2) throw std::runtime_error{"Can't downsample depth texture with 1 mip level"};   // :37-39
  auto depth_ext = depth_desc.extent2D();
  auto norm_ext = norm_desc.extent2D();
  auto vel_ext = vel_desc.extent2D();
  depth_ext.width = std::max(1u, depth_ext.width / 2);
  depth_ext.height = std::max(1u, depth_ext.height / 2);
  if (depth_ext.width != norm_ext.width || depth_ext.height != norm_ext.height || vel_ext.width != norm_ext.width || vel_ext.height != norm_ext.height)
    throw std::runtime_error{"Output textures have different sizes"};                                          // :48-50

  struct Input { ImageViewId gbuffer_depth, gbuffer_normal, gbuffer_velocity, out_depth, out_normal, out_velocity; };
  const bool whole_pyramid = fused;
  graph.add_task<Input>(whole_pyramid ? "DownsampleGbuffer+DownsampleDepth" : "DownsampleGbuffer",
    [&](Input& input, RenderGraphBuilder& builder) {
      input.gbuffer_depth = builder.sample_image(depth, 0, 1, 0, 1);
      input.gbuffer_normal = builder.sample_image(src_normals, 0, 1, 0, 1);
      input.gbuffer_velocity = builder.sample_image(src_velocity, 0, 1, 0, 1);
      input.out_depth = builder.use_depth_attachment(depth, 1, 0);
      input.out_normal = builder.use_color_attachment(out_normal, 0, 0);
      input.out_velocity = builder.use_color_attachment(out_velocity, 0, 0);
    },
    [=](Input& input, RenderResources& resources, gpu::CmdContext& cmd) {
      auto fn = whole_pyramid ? vkr_hiz_build : vkr_downsample_gbuffer;
      vkr_check(fn(resources.get_image(input.gbuffer_depth), resources.get_image(input.gbuffer_normal), resources.get_image(input.gbuffer_velocity),
                   resources.get_image(input.out_normal), resources.get_image(input.out_velocity), cmd.stream), "DownsampleGbuffer");
    });
}

void DownsamplePass::run_downsample_depth(RenderGraph& graph, ImageResourceId depth, uint32_t src_mip) {
  auto desc = graph.get_descriptor(depth);
  struct Input { ImageViewId depth_tex, depth_rt; };
  // The reference adds one task per level (:107-130); one C-ABI call builds them all.
  graph.add_task<Input>("DownsampleDepth",
    [&](Input& input, RenderGraphBuilder& builder) {
      input.depth_rt = builder.use_depth_attachment(depth, desc.mip_levels - 1, 0);
      input.depth_tex = builder.sample_image(depth, src_mip, 1, 0, 1);
    },
    [=](Input& input, RenderResources& resources, gpu::CmdContext& cmd) {
      vkr_check(vkr_depth_mips(resources.get_image(input.depth_tex), src_mip, cmd.stream), "DownsampleDepth");
    });
}

void DownsamplePass::run(RenderGraph& graph, ImageResourceId src_normals, ImageResourceId src_velocity, ImageResourceId depth,
                         ImageResourceId out_normals, ImageResourceId out_velocity) {
  run_downsample_gbuff(graph, src_normals, src_velocity, depth, out_normals, out_velocity);
  if (!fused) run_downsample_depth(graph, depth, 1);                                                            // :141-142
}

// ---- GTAO (reference src/gtao.cpp) ---------------------------------------------------------------
GTAO::GTAO(RenderGraph& graph, uint32_t width, uint32_t height, bool use_ray_query, bool half_res, int pattern_n) : deinterleave_n{pattern_n} {
  if (use_ray_query) throw std::runtime_error{"GTAO: the ray-query variant needs a BVH and is out of scope"};
  if (half_res) { width /= 2; height /= 2; depth_lod = 1; }                                                     // :20-24
  auto mk = [&](uint32_t fmt) { gpu::ImageInfo i; i.format = fmt; i.width = width; i.height = height; return graph.create_image(i); };
  raw = mk(VKR_FORMAT_R16G16B16A16_SFLOAT);                                                                     // :26-33
  filtered = mk(VKR_FORMAT_R16_SFLOAT);
  prev_frame = mk(VKR_FORMAT_R16_SFLOAT);
  output = mk(VKR_FORMAT_R16_SFLOAT);
  accumulated_ao = mk(VKR_FORMAT_R16G16_SFLOAT);                                                                // :37-39
  accumulated_history = mk(VKR_FORMAT_R16G16_SFLOAT);
  if (pattern_n < 0 || pattern_n > 3) throw std::runtime_error{"GTAO: pattern_n outside 0..3"};
  const uint32_t pattern_step = 1u << (uint32_t)pattern_n;                                                      // :41-47
  gpu::ImageInfo di; di.format = VKR_FORMAT_R32_SFLOAT; di.width = width / pattern_step; di.height = height / pattern_step;
  di.array_layers = pattern_step * pattern_step;
  if (di.width && di.height) deinterleaved_depth = graph.create_image(di);
}

void GTAO::add_main_pass(RenderGraph& graph, const GTAOParams& params, ImageResourceId depth, ImageResourceId normal, ImageResourceId material,
                         ImageResourceId preintegrated_pdf) {
  struct PassData { ImageViewId out, depth, norm, material, pdf; };
  const float angle_offsets[]{60.f, 300.f, 180.f, 240.f, 120.f, 0.f, 300.f, 60.f, 180.f, 120.f, 240.f, 0.f};     // :109
  float base_angle = angle_offsets[frame_count % (sizeof(angle_offsets) / sizeof(float))] / 360.f;
  base_angle += angle_jitter;                                                                                   // :111 (explicit input)
  last_angle_offset = base_angle;
  VkrGtaoParams p{};
  copy_mat(p.normal_mat, params.normal_mat);
  p.fovy = params.fovy; p.aspect = params.aspect; p.znear = params.znear; p.zfar = params.zfar;
  p.angle_offset = base_angle; p.weight_ratio = weight_ratio; p.use_mis = mis_gtao;
  p.two_directions = two_directions ? 255u : 0u; p.reflections_only = only_reflections ? 255u : 0u;            // :113
  frame_count += 1;
  const uint32_t lod = depth_lod;
  graph.add_task<PassData>("GTAO_main",
    [&](PassData& input, RenderGraphBuilder& builder) {
      input.depth = builder.sample_image(depth, lod, 1, 0, 1);
      input.norm = builder.sample_image(normal);
      input.material = builder.sample_image(material);
      input.pdf = builder.sample_image(preintegrated_pdf);
      input.out = builder.use_storage_image(raw, 0, 0);
    },
    [=](PassData& input, RenderResources& resources, gpu::CmdContext& cmd) {
      vkr_check(vkr_gtao_main(&p, resources.get_image(input.depth), resources.get_image(input.norm), resources.get_image(input.material),
                              resources.get_image(input.pdf), resources.get_image(input.out), lod, cmd.stream), "GTAO_main");
    });
}

void GTAO::add_filter_pass(RenderGraph& graph, const GTAOParams& params, ImageResourceId depth) {
  struct PassData { ImageViewId out, depth, raw_gtao; };
  VkrGtaoParams p{};
  p.znear = params.znear; p.zfar = params.zfar;                                                                 // :211-216
  const uint32_t lod = depth_lod;
  graph.add_task<PassData>("GTAO_filter",
    [&](PassData& input, RenderGraphBuilder& builder) {
      input.depth = builder.sample_image(depth, lod, 1, 0, 1);
      input.raw_gtao = builder.sample_image(raw);
      input.out = builder.use_storage_image(filtered, 0, 0);
    },
    [=](PassData& input, RenderResources& resources, gpu::CmdContext& cmd) {
      vkr_check(vkr_gtao_filter(&p, resources.get_image(input.depth), resources.get_image(input.raw_gtao), resources.get_image(input.out), lod, cmd.stream),
                "GTAO_filter");
    });
}

void GTAO::add_accumulate_pass(RenderGraph& graph, const DrawTAAParams& params, const Gbuffer& gbuffer) {
  struct PassData { ImageViewId depth, prev_depth, gtao, accumulated_ao, velocity, history; };
  VkrAccumParams p{};
  copy_mat(p.inverse_camera, glm::inverse(params.camera));                                                      // :311
  copy_mat(p.prev_inverse_camera, glm::inverse(params.prev_camera));
  copy_mat(p.mvp, params.mvp);
  std::memcpy(p.fovy_aspect_znear_zfar, &params.fovy_aspect_znear_zfar, 16);
  p.clear_history = clear_history ? 1u : 0u;
  clear_history = false;
  const uint32_t lod = depth_lod;
  graph.add_task<PassData>("GTAO_accumulate",
    [&](PassData& input, RenderGraphBuilder& builder) {
      input.depth = builder.sample_image(gbuffer.depth, lod, 1, 0, 1);
      input.prev_depth = builder.sample_image(gbuffer.prev_depth, lod, 1, 0, 1);
      input.gtao = builder.sample_image(filtered);
      input.accumulated_ao = builder.use_storage_image(accumulated_ao, 0, 0);
      input.velocity = builder.sample_image(gbuffer.downsampled_velocity_vectors);
      input.history = builder.sample_image(accumulated_history);
    },
    [=](PassData& input, RenderResources& resources, gpu::CmdContext& cmd) {
      vkr_check(vkr_gtao_accumulate(&p, resources.get_image(input.depth), resources.get_image(input.prev_depth), resources.get_image(input.gtao),
                                    resources.get_image(input.accumulated_ao), resources.get_image(input.velocity), resources.get_image(input.history),
                                    lod, cmd.stream), "GTAO_accumulate");
    });
}

// ---- GTAO variants beside the live path (reference src/gtao.cpp:241-284,:349-413,:445-526) ----------------------
static const float kGtaoAngles[]{60.f, 300.f, 180.f, 240.f, 120.f, 0.f, 300.f, 60.f, 180.f, 120.f, 240.f, 0.f};

void GTAO::add_main_pass_graphics(RenderGraph& graph, const GTAOParams& params, ImageResourceId depth, ImageResourceId normal) {
  struct PassData { ImageViewId rt, depth, norm; };
  float base_angle = kGtaoAngles[frame_count % (sizeof(kGtaoAngles) / sizeof(float))] / 360.f;                   // :370-372
  base_angle += angle_jitter;
  frame_count += 1;
  last_angle_offset = base_angle;
  VkrGtaoParams p{};
  copy_mat(p.normal_mat, params.normal_mat);
  p.fovy = params.fovy; p.aspect = params.aspect; p.znear = params.znear; p.zfar = params.zfar;
  p.angle_offset = base_angle;
  const uint32_t lod = depth_lod;
  graph.add_task<PassData>("GTAO",
    [&](PassData& input, RenderGraphBuilder& builder) {
      input.depth = builder.sample_image(depth, depth_lod, 1, 0, 1);
      input.norm = builder.sample_image(normal);
      input.rt = builder.use_color_attachment(raw, 0, 0);
    },
    [=](PassData& input, RenderResources& resources, gpu::CmdContext& cmd) {
      vkr_check(vkr_gtao_main_graphics(&p, resources.get_image(input.depth), resources.get_image(input.norm), resources.get_image(input.rt), lod, nullptr,
                                       cmd.stream), "GTAO");
    });
}

void GTAO::add_reprojection_pass(RenderGraph& graph, const GTAOReprojection& params, ImageResourceId depth, ImageResourceId prev_depth) {
  struct PassData { ImageViewId out, gtao, prev_gtao, depth, prev_depth; };
  struct Pc { float v[4]; } pc{{params.fovy, params.aspect, params.znear, params.zfar}};
  const uint32_t lod = depth_lod;
  graph.add_task<PassData>("GTAO_reproject",
    [&](PassData& input, RenderGraphBuilder& builder) {
      input.depth = builder.sample_image(depth, depth_lod, 1, 0, 1);
      input.prev_depth = builder.sample_image(prev_depth, depth_lod, 1, 0, 1);
      input.gtao = builder.sample_image(filtered);
      input.prev_gtao = builder.sample_image(prev_frame);
      input.out = builder.use_storage_image(output, 0, 0);
    },
    [=](PassData& input, RenderResources& resources, gpu::CmdContext& cmd) {
      vkr_check(vkr_gtao_reproject(pc.v, resources.get_image(input.depth), resources.get_image(input.prev_depth), lod, resources.get_image(input.gtao),
                                   resources.get_image(input.prev_gtao), resources.get_image(input.out), nullptr, cmd.stream), "GTAO_reproject");
    });
}

void GTAO::deinterleave_depth(RenderGraph& graph, ImageResourceId depth) {
  struct PassData { ImageViewId depth, out; };
  const uint32_t lod = depth_lod;
  const int n = deinterleave_n;
  const uint64_t stride = graph.layer_stride(deinterleaved_depth);
  graph.add_task<PassData>("GTAO_deinterleave",
    [&](PassData& input, RenderGraphBuilder& builder) {
      input.depth = builder.sample_image(depth, depth_lod, 1, 0, 1);
      input.out = builder.use_storage_image(deinterleaved_depth, 0, 0);
    },
    [=](PassData& input, RenderResources& resources, gpu::CmdContext& cmd) {
      vkr_check(vkr_gtao_deinterleave(resources.get_image(input.depth), lod, resources.get_image(input.out), stride, n, cmd.stream), "GTAO_deinterleave");
    });
}

void GTAO::add_main_pass_deinterleaved(RenderGraph& graph, const GTAOParams& params, ImageResourceId normal) {
  struct PassData { ImageViewId out, depth, norm; };
  float base_angle = kGtaoAngles[frame_count % (sizeof(kGtaoAngles) / sizeof(float))] / 360.f;                   // :493-495
  base_angle += angle_jitter;
  frame_count += 1;
  last_angle_offset = base_angle;
  VkrGtaoParams p{};
  copy_mat(p.normal_mat, params.normal_mat);
  p.fovy = params.fovy; p.aspect = params.aspect; p.znear = params.znear; p.zfar = params.zfar;
  p.angle_offset = base_angle;
  const int n = deinterleave_n;
  const uint64_t stride = graph.layer_stride(deinterleaved_depth);
  const uint32_t out_layers = graph.get_descriptor(raw).array_layers;                                            // :520 — the OUTPUT image's layers (1)
  graph.add_task<PassData>("GTAO_deinterleaved",
    [&](PassData& input, RenderGraphBuilder& builder) {
      input.depth = builder.sample_image(deinterleaved_depth);
      input.norm = builder.sample_image(normal);
      input.out = builder.use_storage_image(raw, 0, 0);
    },
    [=](PassData& input, RenderResources& resources, gpu::CmdContext& cmd) {
      for (uint32_t i = 0; i < out_layers; i++)
        vkr_check(vkr_gtao_main_deinterleaved(&p, resources.get_image(input.depth), stride, resources.get_image(input.norm), resources.get_image(input.out), n, i,
                                              nullptr, cmd.stream), "GTAO_deinterleaved");
    });
}
