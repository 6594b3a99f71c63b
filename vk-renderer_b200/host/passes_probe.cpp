// passes_probe.cpp — OctahedralProbe(Grid), ProbeRenderer, ProbeTracePass task bodies (see passes.hpp; reference src/probe_renderer.cpp).
#include <algorithm>
#include <cmath>

#include "../../include/vkr_probe.h"
#include "../../include/vkr_synth.h"
#include "passes.hpp"

using namespace rendergraph;

OctahedralProbe::OctahedralProbe(RenderGraph& graph, uint32_t size) {                                            // :285-295
  gpu::ImageInfo desc; desc.format = VKR_FORMAT_R8G8B8A8_UNORM; desc.width = size; desc.height = size;
  color = graph.create_image(desc);
  desc.mip_levels = (uint32_t)std::floor(std::log2((double)size)) + 1;
  desc.format = VKR_FORMAT_R16_UNORM;
  depth = graph.create_image(desc);
}

OctahedralProbeGrid::OctahedralProbeGrid(RenderGraph& graph, uint32_t grid_sz, uint32_t size) : grid_size{grid_sz} {   // :297-314
  gpu::ImageInfo desc; desc.format = VKR_FORMAT_R8G8B8A8_UNORM; desc.width = size; desc.height = size;
  desc.array_layers = grid_sz * grid_sz;
  color_array = graph.create_image(desc);
  desc.mip_levels = (uint32_t)std::floor(std::log2((double)size)) + 1;
  desc.format = VKR_FORMAT_R16_UNORM;
  depth_array = graph.create_image(desc);
}

ProbeRenderer::ProbeRenderer(RenderGraph& graph, uint32_t cubemap_res) {                                          // :9-24
  gpu::ImageInfo desc; desc.format = VKR_FORMAT_R8G8B8A8_SRGB; desc.width = cubemap_res; desc.height = cubemap_res; desc.array_layers = 6;
  cubemap_color = graph.create_image(desc);
  desc.format = VKR_FORMAT_R16_SFLOAT;
  cubemap_distance = graph.create_image(desc);
}

void ProbeRenderer::render_cubemap(RenderGraph& graph, const VkrSynthScene& scene, const vec3 pos) {
  struct Input { ImageViewId color, distance; };
  const uint64_t cs = graph.layer_stride(cubemap_color), ds = graph.layer_stride(cubemap_distance);
  const VkrSynthScene sc = scene;
  graph.add_task<Input>("CubemapProbe",
    [&](Input& input, RenderGraphBuilder& builder) {
      input.color = builder.use_color_attachment(cubemap_color, 0, 0);
      input.distance = builder.use_color_attachment(cubemap_distance, 0, 0);
    },
    [=](Input& input, RenderResources& resources, gpu::CmdContext& cmd) {
      const float p[3] = {pos.x, pos.y, pos.z};
      vkr_check(vkr_synth_cubemap(&sc, p, resources.get_image(input.color), cs, resources.get_image(input.distance), ds, cmd.stream), "CubemapProbe");
    });
}

void ProbeRenderer::render_octahedral(RenderGraph& graph, ImageResourceId probe_color, ImageResourceId probe_depth, uint32_t array_layer) {
  struct Input { ImageViewId cube_color, cube_distance, oct_color, oct_depth; };
  const uint64_t cs = graph.layer_stride(cubemap_color), ds = graph.layer_stride(cubemap_distance);
  const uint64_t ocs = graph.layer_stride(probe_color), ods = graph.layer_stride(probe_depth);
  graph.add_task<Input>("Cubemap2Octahedral",                                                                      // :179
    [&](Input& input, RenderGraphBuilder& builder) {
      input.cube_color = builder.sample_image(cubemap_color);
      input.cube_distance = builder.sample_image(cubemap_distance);
      input.oct_color = builder.use_storage_image(probe_color, 0, array_layer);
      input.oct_depth = builder.use_storage_image(probe_depth, 0, array_layer);
    },
    [=](Input& input, RenderResources& resources, gpu::CmdContext& cmd) {
      vkr_check(vkr_cube2oct(resources.get_image(input.cube_color), cs, resources.get_image(input.cube_distance), ds, resources.get_image(input.oct_color), ocs,
                             resources.get_image(input.oct_depth), ods, array_layer, 1, cmd.stream), "Cubemap2Octahedral");
    });
}

void ProbeRenderer::probe_downsample(RenderGraph& graph, ImageResourceId probe_depth, uint32_t array_layer) {
  struct Input { ImageViewId depth_tex; };
  const uint64_t ods = graph.layer_stride(probe_depth);
  // the reference adds one "DownsampleProbe" task per mip (:215-235); the C entry walks the mips of the layer in one call
  graph.add_task<Input>("DownsampleProbe",
    [&](Input& input, RenderGraphBuilder& builder) { input.depth_tex = builder.use_color_attachment(probe_depth, 1, array_layer); },
    [=](Input& input, RenderResources& resources, gpu::CmdContext& cmd) {
      vkr_check(vkr_probe_downsample(resources.get_image(input.depth_tex), ods, array_layer, 1, cmd.stream), "DownsampleProbe");
    });
}

void ProbeRenderer::render_probe(RenderGraph& graph, const VkrSynthScene& scene, const vec3 pos, OctahedralProbe& probe) {   // :238-243
  render_cubemap(graph, scene, pos);
  render_octahedral(graph, probe.color, probe.depth);
  probe_downsample(graph, probe.depth);
  probe.pos = pos;
}

static void swap_min(float& mn, float& mx) { float a = std::min(mn, mx), b = std::max(mn, mx); mn = a; mx = b; }

void ProbeRenderer::render_probe_grid(RenderGraph& graph, const VkrSynthScene& scene, vec3 min, vec3 max, OctahedralProbeGrid& probe_grid) {   // :253-283
  swap_min(min.x, max.x);
  swap_min(min.y, max.y);
  swap_min(min.z, max.z);
  probe_grid.min = min;
  probe_grid.max = max;
  if (probe_grid.grid_size < 2) throw std::runtime_error{"Ooops"};
  const float denom = float(probe_grid.grid_size - 1);
  const vec3 step{(max.x - min.x) / denom, (max.y - min.y) / denom, (max.z - min.z) / denom};
  for (uint32_t y = 0; y < probe_grid.grid_size; y++)
    for (uint32_t x = 0; x < probe_grid.grid_size; x++) {
      const vec3 pos{min.x + step.x * float(x), min.y + step.y * 0.f, min.z + step.z * float(y)};
      const uint32_t array_layer = y * probe_grid.grid_size + x;
      render_cubemap(graph, scene, pos);
      render_octahedral(graph, probe_grid.color_array, probe_grid.depth_array, array_layer);
      probe_downsample(graph, probe_grid.depth_array, array_layer);
    }
}

void ProbeTracePass::run(RenderGraph& graph, OctahedralProbeGrid& probe, ImageResourceId gbuffer_depth, ImageResourceId gbuffer_norm, ImageResourceId out_image,
                         const ProbeTraceParams& params) {
  struct Input { ImageViewId depth, normal, probe_color, probe_depth, out_tex; };
  VkrProbeTraceParams consts{};                                                                                    // :340-360
  std::memcpy(consts.inverse_view, params.inv_view.m, 64);
  consts.probe_min[0] = probe.min.x; consts.probe_min[1] = probe.min.y; consts.probe_min[2] = probe.min.z; consts.probe_min[3] = 1.f;
  consts.probe_max[0] = probe.max.x; consts.probe_max[1] = probe.max.y; consts.probe_max[2] = probe.max.z; consts.probe_max[3] = 1.f;
  consts.grid_size = probe.grid_size;
  consts.fovy = params.fovy; consts.aspect = params.aspect; consts.znear = params.znear; consts.zfar = params.zfar;
  const uint64_t cs = graph.layer_stride(probe.color_array), ds = graph.layer_stride(probe.depth_array);
  graph.add_task<Input>("TraceProbe",
    [&](Input& input, RenderGraphBuilder& builder) {
      input.depth = builder.sample_image(gbuffer_depth, 0, 1, 0, 1);
      input.normal = builder.sample_image(gbuffer_norm);
      input.probe_color = builder.sample_image(probe.color_array);
      input.probe_depth = builder.sample_image(probe.depth_array);
      input.out_tex = builder.use_storage_image(out_image, 0, 0);
    },
    [=](Input& input, RenderResources& resources, gpu::CmdContext& cmd) {
      vkr_check(vkr_trace_probe(&consts, resources.get_image(input.depth), resources.get_image(input.normal), resources.get_image(input.probe_color), cs,
                                resources.get_image(input.probe_depth), ds, resources.get_image(input.out_tex), nullptr, cmd.stream), "TraceProbe");
    });
}
