// passes.hpp — the reference's pass classes with the same constructors, methods and public image
// members, their task bodies calling the CUDA C ABI (include/vkr_passes.h) instead of recording
// Vulkan commands.  Reference surfaces mirrored:
//   struct Gbuffer, DrawTAAParams            src/scene_renderer.hpp:11-33, src/scene_renderer.cpp:8-44
//   DownsamplePass                           src/downsample_pass.hpp:6-36
//   GTAO, GTAOParams                         src/gtao.hpp:12-121
//   AdvancedSSR, AdvancedSSRParams           src/advanced_ssr.hpp:7-116
//   DeferedShadingPass                       src/defered_shading.hpp:8-32
//   TAA                                      src/taa.hpp:8-21
//   SSAOPass, SSAOInParams, create_ssao_texture   src/ssao.hpp:7-31
//   SSRParams, add_ssr_pass, create_ssr_tex       src/ssr.hpp:7-26
//   ScreenSpaceTrace, ScreenTraceParams            src/screen_trace.hpp:8-51
// Dropped: pipelines / samplers / descriptor sets (no graphics API), the SDL_Window* of
// DeferedShadingPass (ImGui only), ray-query and graphics-pipeline GTAO variants, ImGui panels.
#pragma once
#include <cstring>
#include <vector>

#include "../../include/vkr_synth.h"
#include "rendergraph.hpp"

namespace glmlite {  // the two glm types the pass interface exposes
struct mat4 {
  float m[16];  // column-major, as glm::mat4
  mat4() { std::memset(m, 0, sizeof(m)); m[0] = m[5] = m[10] = m[15] = 1.f; }
  explicit mat4(const float* p) { std::memcpy(m, p, sizeof(m)); }
};
struct vec4 { float x = 0, y = 0, z = 0, w = 0; };
struct vec2 { float x = 0, y = 0; };
mat4 inverse(const mat4& a);
mat4 transpose(const mat4& a);
}  // namespace glmlite
namespace glm = glmlite;

struct Gbuffer {
  Gbuffer(rendergraph::RenderGraph& graph, uint32_t width, uint32_t height);
  rendergraph::ImageResourceId albedo;
  rendergraph::ImageResourceId normal;
  rendergraph::ImageResourceId downsampled_normals;
  rendergraph::ImageResourceId material;
  rendergraph::ImageResourceId depth;
  rendergraph::ImageResourceId prev_depth;
  rendergraph::ImageResourceId velocity_vectors;
  rendergraph::ImageResourceId downsampled_velocity_vectors;
  uint32_t w, h;
};

struct DrawTAAParams {
  glm::mat4 mvp;
  glm::mat4 prev_mvp;
  glm::mat4 camera;
  glm::mat4 prev_camera;
  glm::vec4 jitter;
  glm::vec4 fovy_aspect_znear_zfar;
};

struct DownsamplePass {
  DownsamplePass() {}
  // fused = one launch for the half-res G-buffer and pyramid levels 1..6 plus one tail launch;
  // unfused issues the reference's task list ("DownsampleGbuffer", "DownsampleDepth").
  bool fused = true;
  void run(rendergraph::RenderGraph& graph, rendergraph::ImageResourceId src_normals, rendergraph::ImageResourceId src_velocity,
           rendergraph::ImageResourceId depth, rendergraph::ImageResourceId out_normals, rendergraph::ImageResourceId out_velocity);

 private:
  void run_downsample_gbuff(rendergraph::RenderGraph& graph, rendergraph::ImageResourceId src_normals, rendergraph::ImageResourceId src_velocity,
                            rendergraph::ImageResourceId depth, rendergraph::ImageResourceId out_normal, rendergraph::ImageResourceId out_velocity);
  void run_downsample_depth(rendergraph::RenderGraph& graph, rendergraph::ImageResourceId depth, uint32_t src_mip);
};

struct GTAOParams {
  glm::mat4 normal_mat;
  float fovy;
  float aspect;
  float znear;
  float zfar;
};

struct GTAOReprojection {
  glm::mat4 camera_to_prev_frame;   // read only by the MATRIX_REPROJECT mode the reference compiles out (reproject.comp:4-7)
  float fovy;
  float aspect;
  float znear;
  float zfar;
};

struct GTAO {
  GTAO(rendergraph::RenderGraph& graph, uint32_t width, uint32_t height, bool use_ray_query, bool half_res = true, int pattern_n = 2);

  void add_main_pass(rendergraph::RenderGraph& graph, const GTAOParams& params, rendergraph::ImageResourceId depth,
                     rendergraph::ImageResourceId normal, rendergraph::ImageResourceId material, rendergraph::ImageResourceId preintegrated_pdf);
  void add_filter_pass(rendergraph::RenderGraph& graph, const GTAOParams& params, rendergraph::ImageResourceId depth);
  void add_accumulate_pass(rendergraph::RenderGraph& graph, const DrawTAAParams& params, const Gbuffer& gbuffer);
  // the variants beside the live path (reference src/gtao.hpp:57-82); each advances the same frame_count / angle table
  void add_main_pass_graphics(rendergraph::RenderGraph& graph, const GTAOParams& params, rendergraph::ImageResourceId depth,
                              rendergraph::ImageResourceId normal);
  void add_reprojection_pass(rendergraph::RenderGraph& graph, const GTAOReprojection& params, rendergraph::ImageResourceId depth,
                             rendergraph::ImageResourceId prev_depth);
  void deinterleave_depth(rendergraph::RenderGraph& graph, rendergraph::ImageResourceId depth);
  void add_main_pass_deinterleaved(rendergraph::RenderGraph& graph, const GTAOParams& params, rendergraph::ImageResourceId normal);
  void remap(rendergraph::RenderGraph& graph) { graph.remap(accumulated_history, accumulated_ao); }

  rendergraph::ImageResourceId raw;       // output of main pass
  rendergraph::ImageResourceId filtered;  // output of filter pass
  rendergraph::ImageResourceId prev_frame;
  rendergraph::ImageResourceId output;
  rendergraph::ImageResourceId accumulated_ao;
  rendergraph::ImageResourceId accumulated_history;
  rendergraph::ImageResourceId deinterleaved_depth;   // R32F, (w >> n) x (h >> n), 4^n layers (gtao.cpp:41-47)
  float last_angle_offset = 0.f;                       // push constant of the most recent main-pass variant

  // runtime settings (ImGui state in the reference, src/gtao.hpp:108-116)
  bool mis_gtao = true;
  bool two_directions = false;
  bool only_reflections = false;
  bool clear_history = false;
  float weight_ratio = 1.0;
  // The reference adds rand()/RAND_MAX - 0.5 to the per-frame angle (src/gtao.cpp:111); here the
  // random part is an explicit input so runs are reproducible.
  float angle_jitter = 0.f;

 private:
  int deinterleave_n = 2;
  uint32_t depth_lod = 0;
  uint32_t frame_count = 0;
};

struct AdvancedSSRParams {
  glm::mat4 normal_mat;
  float fovy;
  float aspect;
  float znear;
  float zfar;
};

std::vector<glm::vec4> halton23_seq(uint32_t count);

struct AdvancedSSR {
  AdvancedSSR(rendergraph::RenderGraph& graph, uint32_t w, uint32_t h);
  ~AdvancedSSR();
  void run(rendergraph::RenderGraph& graph, const AdvancedSSRParams& params, const DrawTAAParams& taa_params, const Gbuffer& gbuff,
           rendergraph::ImageResourceId ssr_occlusion);
  void preintegrate_pdf(rendergraph::RenderGraph& graph);
  void preintegrate_brdf(rendergraph::RenderGraph& graph);
  void remap_images(rendergraph::RenderGraph& graph) { graph.remap(blurred_reflection, blurred_reflection_history); }

  rendergraph::ImageResourceId get_ouput() const { return reflections; }
  rendergraph::ImageResourceId get_rays() const { return rays; }
  rendergraph::ImageResourceId get_blurred() const { return blurred_reflection; }
  rendergraph::ImageResourceId get_blurred_history() const { return blurred_reflection_history; }  // not in the reference: for readback after remap
  rendergraph::ImageResourceId get_preintegrated_pdf() const { return preintegrated_pdf; }
  rendergraph::ImageResourceId get_preintegrated_brdf() const { return preintegrated_brdf; }

  struct {
    float max_rougness = 1.f;
    float glossy_roughness_value = 0.5f;
    bool normalize_reflections = true;
    bool accumulate_reflections = true;
    bool bilateral_filter = true;
    bool update_random = true;
    bool use_blur = true;
    int max_accumulated_rays = 16;
    // not in the reference: run() issues the four tile tasks the reference has commented out at src/advanced_ssr.cpp:547-550
    // (clear, classification, tile regression, indirect trace) ahead of the per-pixel trace, in that order
    bool run_tile_path = false;
  } settings;

  // The tile path (reference src/advanced_ssr.cpp:216-302,:440-538; private there and unreachable from run()).  Public here so
  // a host can schedule the tasks itself.
  void clear_indirect_params(rendergraph::RenderGraph& graph);
  void run_classification_pass(rendergraph::RenderGraph& graph, const AdvancedSSRParams& params, const Gbuffer& gbuff);
  void run_tile_regression_pass(rendergraph::RenderGraph& graph, const AdvancedSSRParams& params, const Gbuffer& gbuff);
  void run_trace_indirect_pass(rendergraph::RenderGraph& graph, const AdvancedSSRParams& params, const Gbuffer& gbuff);
  rendergraph::ImageResourceId get_tile_planes() const { return tile_planes; }
  rendergraph::BufferResourceId get_reflective_indirect() const { return reflective_indirect; }
  rendergraph::BufferResourceId get_glossy_indirect() const { return glossy_indirect; }
  rendergraph::BufferResourceId get_reflective_tiles() const { return reflective_tiles; }
  rendergraph::BufferResourceId get_glossy_tiles() const { return glossy_tiles; }

 private:
  rendergraph::BufferResourceId reflective_indirect;
  rendergraph::BufferResourceId glossy_indirect;
  rendergraph::BufferResourceId reflective_tiles;
  rendergraph::BufferResourceId glossy_tiles;
  rendergraph::ImageResourceId tile_planes;
  float* halton_buffer = nullptr;  // device, 128 x vec4 (the Halton UBO)
  rendergraph::ImageResourceId rays;
  rendergraph::ImageResourceId reflections;
  rendergraph::ImageResourceId blurred_reflection;
  rendergraph::ImageResourceId blurred_reflection_history;
  rendergraph::ImageResourceId preintegrated_pdf;
  rendergraph::ImageResourceId preintegrated_brdf;
  uint32_t counter{0u};

  void run_trace_pass(rendergraph::RenderGraph& graph, const AdvancedSSRParams& params, const Gbuffer& gbuff, rendergraph::ImageResourceId ssr_occlusion);
  void run_filter_pass(rendergraph::RenderGraph& graph, const AdvancedSSRParams& params, const Gbuffer& gbuff);
  void run_blur_pass(rendergraph::RenderGraph& graph, const AdvancedSSRParams& params, const DrawTAAParams& taa_params, const Gbuffer& gbuff);
};

struct DeferedShadingPass {
  explicit DeferedShadingPass(rendergraph::RenderGraph& graph);
  void update_params(const glm::mat4& camera, const glm::mat4& shadow, float fovy, float aspect, float znear, float zfar);
  // `shadow` is accepted like in the reference and, like there, never read by the shading code.
  void draw(rendergraph::RenderGraph& graph, const Gbuffer& gbuffer, rendergraph::ImageResourceId shadow, rendergraph::ImageResourceId ssao,
            rendergraph::ImageResourceId brdf_tex, rendergraph::ImageResourceId reflections, rendergraph::ImageResourceId out_image);

  glm::vec2 min_max_roughness{0.f, 1.f};
  bool only_ao = false;

 private:
  VkrShadingParams consts{};
};

struct TAA {
  TAA(rendergraph::RenderGraph& graph, uint32_t w, uint32_t h);
  void run(rendergraph::RenderGraph& graph, const Gbuffer& gbuffer, rendergraph::ImageResourceId color, const DrawTAAParams& params);
  void remap_targets(rendergraph::RenderGraph& graph) { graph.remap(history, target); }
  rendergraph::ImageResourceId get_output() const { return target; }
  rendergraph::ImageResourceId get_history() const { return history; }  // not in the reference: for readback after remap

 private:
  rendergraph::ImageResourceId history;
  rendergraph::ImageResourceId target;
};

// ---- the earlier passes (present in the reference tree, not scheduled by its frame loop) -------------------------
rendergraph::ImageResourceId create_ssao_texture(rendergraph::RenderGraph& graph, uint32_t width, uint32_t height);

struct SSAOInParams {
  glm::mat4 projection;
  float fovy;
  float aspect;
  float znear;
  float zfar;
};

struct vec3 { float x = 0, y = 0, z = 0; };

struct SSAOPass {
  // The reference seeds nothing and draws from libc rand() (src/ssao.cpp:34-52); `seed` makes the sample set an input.
  SSAOPass(rendergraph::RenderGraph& graph, rendergraph::ImageResourceId target, uint32_t seed = 0);
  void draw(rendergraph::RenderGraph& graph, rendergraph::ImageResourceId depth, rendergraph::ImageResourceId target, const SSAOInParams& params);

  std::vector<vec3> sphere_samples;   // private in the reference; public here so tests can hand the same set to the oracle
  // true: upload the block with the reference's C++ layout (vec3 at 12 bytes) and let the kernel read it at the shader's
  // 16-byte stride, like the reference does; false: each sample in its own std140 slot.
  bool reference_packing = false;
  VkrSsaoParams last_block{};         // the block of the most recent draw() (what the reference writes into its UBO)
};

rendergraph::ImageResourceId create_ssr_tex(rendergraph::RenderGraph& graph, uint32_t w, uint32_t h);

struct SSRParams {
  glm::mat4 normal_mat;
  float fovy;
  float aspect;
  float znear;
  float zfar;
};

void add_ssr_pass(rendergraph::RenderGraph& graph, rendergraph::ImageResourceId depth, rendergraph::ImageResourceId normal,
                  rendergraph::ImageResourceId color, rendergraph::ImageResourceId material, rendergraph::ImageResourceId out, const SSRParams& params);

struct ScreenTraceParams {
  glm::mat4 normal_mat;
  float fovy;
  float aspect;
  float znear;
  float zfar;
};

struct ScreenSpaceTrace {
  ScreenSpaceTrace(rendergraph::RenderGraph& graph, uint32_t width, uint32_t height);

  void add_main_pass(rendergraph::RenderGraph& graph, const ScreenTraceParams& params, rendergraph::ImageResourceId depth,
                     rendergraph::ImageResourceId normal, rendergraph::ImageResourceId color, rendergraph::ImageResourceId material);
  void add_filter_pass(rendergraph::RenderGraph& graph, const ScreenTraceParams& params, rendergraph::ImageResourceId depth);
  void add_accumulate_pass(rendergraph::RenderGraph& graph, const ScreenTraceParams& params, rendergraph::ImageResourceId depth,
                           rendergraph::ImageResourceId prev_depth);

  rendergraph::ImageResourceId raw;
  rendergraph::ImageResourceId filtered;
  rendergraph::ImageResourceId accumulated;

  // The reference draws both from std::default_random_engine inside add_main_pass (src/screen_trace.cpp:49-54); here the
  // two uniforms in [0,1) of the next call are inputs so runs are reproducible.
  float next_angle_random = 0.5f, next_random_offset = 0.f;
  VkrScreenTraceParams last_block{};   // the block of the most recent add_main_pass()

 private:
  uint32_t frame_count = 0;
};

// ---- probe_renderer (reference src/probe_renderer.hpp:6-81) ---------------------------------------------------------
const uint32_t PROBE_SIZE = 256;
const uint32_t CUBE_SIZE = 128;

struct OctahedralProbe {
  OctahedralProbe(rendergraph::RenderGraph& graph, uint32_t size = PROBE_SIZE);
  vec3 pos{0.f, 0.f, 0.f};
  rendergraph::ImageResourceId color;
  rendergraph::ImageResourceId depth;
};

struct OctahedralProbeGrid {
  OctahedralProbeGrid(rendergraph::RenderGraph& graph, uint32_t grid_sz = 4, uint32_t size = PROBE_SIZE);
  vec3 min{0.f, 0.f, 0.f};
  vec3 max{0.f, 0.f, 0.f};
  uint32_t grid_size{0};
  rendergraph::ImageResourceId color_array;
  rendergraph::ImageResourceId depth_array;
};

// The reference rasterises the cube faces from its SceneRenderer (render_cubemap / render_side, :78-168); here the faces come
// from the synthetic scene's ray-caster (vkr_synth_cubemap), so `SceneRenderer&` becomes `const VkrSynthScene&`.
struct ProbeRenderer {
  ProbeRenderer(rendergraph::RenderGraph& graph, uint32_t cubemap_res = CUBE_SIZE);
  void render_cubemap(rendergraph::RenderGraph& graph, const VkrSynthScene& scene, const vec3 pos);
  void render_probe(rendergraph::RenderGraph& graph, const VkrSynthScene& scene, const vec3 pos, OctahedralProbe& probe);
  void render_probe_grid(rendergraph::RenderGraph& graph, const VkrSynthScene& scene, vec3 min, vec3 max, OctahedralProbeGrid& probe_grid);
  rendergraph::ImageResourceId get_cubemap_color() const { return cubemap_color; }        // not in the reference: for readback
  rendergraph::ImageResourceId get_cubemap_distance() const { return cubemap_distance; }

 private:
  rendergraph::ImageResourceId cubemap_color;
  rendergraph::ImageResourceId cubemap_distance;
  void render_octahedral(rendergraph::RenderGraph& graph, rendergraph::ImageResourceId probe_color, rendergraph::ImageResourceId probe_depth,
                         uint32_t array_layer = 0);
  void probe_downsample(rendergraph::RenderGraph& graph, rendergraph::ImageResourceId probe_depth, uint32_t array_layer = 0);
};

struct ProbeTraceParams {
  glm::mat4 inv_view;  // camera -> world
  float fovy;
  float aspect;
  float znear;
  float zfar;
};

struct ProbeTracePass {
  ProbeTracePass() {}
  void run(rendergraph::RenderGraph& graph, OctahedralProbeGrid& probe, rendergraph::ImageResourceId gbuffer_depth, rendergraph::ImageResourceId gbuffer_norm,
           rendergraph::ImageResourceId out_image, const ProbeTraceParams& params);
};
