/* ORACLE — TEST INFRASTRUCTURE ONLY.
 *
 * C API of the CPU restatement of vk-renderer's image-space shaders.  Loaded with
 * ctypes by tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
 * --impl reference legs — never by the product path (vk-renderer_b200/).
 *
 * PARITY UNPINNED: the reference has no tests, golden images or known-answer vectors
 *  and its GLSL cannot be executed here (no glslc, no Vulkan
 * implementation).  This restatement is pinned only by the analytic anchors in
 * tests/test_oracle_anchors.py.
 *
 * Every pass takes a row range [y0, y1) of OUTPUT rows (y1 <= 0 means "to the end") so
 * a bounded sample of a large frame can be timed; rows are processed with
 * `#pragma omp parallel for schedule(dynamic)`.
 */
#ifndef ORACLE_H_INCLUDED
#define ORACLE_H_INCLUDED
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct OrcImage {
  void* base;
  uint32_t width, height;
  uint32_t format;     /* VkFormat value */
  uint32_t mip_count;
  uint64_t mip_offset[16];
  uint32_t row_pitch[16];
} OrcImage;

/* UBO TraceParams + push constants of SSSR_trace / SSSR_filter
 * (reference src/advanced_ssr.cpp:138-145,:159-163,:328-338). */
typedef struct OrcTraceParams {
  float normal_mat[16];
  uint32_t frame_random;
  float fovy, aspect, znear, zfar;
  float max_roughness;   /* trace push constant */
  uint32_t render_flags; /* filter push constant: NORMALIZE=1 ACCUMULATE=2 BILATERAL=4 */
} OrcTraceParams;

/* UBO + push constants of SSSR_blur (reference src/advanced_ssr.cpp:388-401). */
typedef struct OrcBlurParams {
  float inverse_camera[16];
  float prev_inverse_camera[16];
  float fovy_aspect_znear_zfar[4];
  float max_roughness;
  uint32_t accumulate;
  uint32_t disable_blur;
} OrcBlurParams;

/* GTAOParams UBO + push constants of GTAO_main / GTAO_filter
 * (reference src/gtao.hpp:12-18, src/gtao.cpp:101-113,:211-216). */
typedef struct OrcGtaoParams {
  float normal_mat[16];
  float fovy, aspect, znear, zfar;
  float angle_offset, weight_ratio;
  uint32_t use_mis, two_directions, reflections_only;
} OrcGtaoParams;

/* AccumConstants + push constant of GTAO_accumulate (reference src/gtao.cpp:300-313). */
typedef struct OrcAccumParams {
  float inverse_camera[16];
  float prev_inverse_camera[16];
  float mvp[16];
  float fovy_aspect_znear_zfar[4];
  uint32_t clear_history;
} OrcAccumParams;

/* ShaderConstants + push constants of DeferedShading (reference src/defered_shading.cpp:4-12,:68-72). */
typedef struct OrcShadingParams {
  float inverse_camera[16];
  float camera[16];
  float shadow_mvp[16];
  float fovy, aspect, znear, zfar;
  float min_max_roughness[2];
  uint32_t show_ao;
} OrcShadingParams;

/* TAAParams (reference src/taa.cpp:24-30). */
typedef struct OrcTaaParams {
  float inverse_camera[16];
  float prev_inverse_camera[16];
  float fovy_aspect_znear_zfar[4];
} OrcTaaParams;

/* SSAOParams as the SHADER declares it (reference shaders/ssao/shader.frag:11-19, std140: vec3 array stride 16). */
typedef struct OrcSsaoParams {
  float projection[16];
  float fovy, aspect, znear, zfar;
  float samples[16][4];
} OrcSsaoParams;

/* SSRParams (reference src/ssr.hpp:11-17). */
typedef struct OrcSsrParams {
  float normal_mat[16];
  float fovy, aspect, znear, zfar;
} OrcSsrParams;

/* GpuParams of ScreenSpaceTrace::add_main_pass (reference src/screen_trace.cpp:31-39). */
typedef struct OrcScreenTraceParams {
  float normal_mat[16];
  float random_offset, angle_offset;
  float fovy, aspect, znear, zfar;
} OrcScreenTraceParams;

/* Constants UBO of ProbeTracePass::run (reference src/probe_renderer.cpp:340-360). */
typedef struct OrcProbeTraceParams {
  float inverse_view[16];
  float probe_min[4];
  float probe_max[4];
  uint32_t grid_size;
  float fovy, aspect, znear, zfar;
} OrcProbeTraceParams;

void orc_set_threads(int n);
int  orc_get_max_threads(void);

/* reference src/advanced_ssr.cpp:8-34 — 128 x vec4 (x = base 2, y = base 3, index+1). */
void orc_halton23_seq(uint32_t count, float* out_vec4);

/* Hi-Z: reference src/downsample_pass.cpp:25-131, shaders/advanced_ssr/{downsample_gbuffer,depth_mips}.frag */
int orc_downsample_gbuffer(const OrcImage* depth, const OrcImage* normal, const OrcImage* velocity,
                           const OrcImage* out_normal, const OrcImage* out_velocity, int y0, int y1);
int orc_depth_mips(const OrcImage* depth, uint32_t src_mip);

/* SSR: reference shaders/advanced_ssr/{trace,filter,blur,preintegrate,preintegrate_ssr}.comp */
int orc_sssr_trace(const OrcTraceParams* p, const OrcImage* depth, const OrcImage* dnormal, const OrcImage* material,
                   const float* halton_vec4x128, const OrcImage* pdf_lut, const OrcImage* out_rays,
                   const OrcImage* out_occlusion, int y0, int y1);
int orc_sssr_filter(const OrcTraceParams* p, const OrcImage* rays, const OrcImage* depth, const OrcImage* albedo,
                    const OrcImage* normal, const OrcImage* material, const OrcImage* out_reflections, int y0, int y1);
int orc_sssr_blur(const OrcBlurParams* p, const OrcImage* depth, const OrcImage* normal, const OrcImage* reflections,
                  const OrcImage* material, const OrcImage* history, const OrcImage* dvelocity,
                  const OrcImage* prev_depth, const OrcImage* out_blurred, int y0, int y1);
int orc_pdf_preintegrate(const OrcImage* out_pdf, int y0, int y1);
int orc_brdf_preintegrate(const float* halton_vec4x128, const OrcImage* out_brdf, int y0, int y1);

/* GTAO: reference shaders/gtao/{main,filter,accum}.comp */
int orc_gtao_main(const OrcGtaoParams* p, const OrcImage* depth, const OrcImage* normal, const OrcImage* material,
                  const OrcImage* pdf_lut, const OrcImage* raw, uint32_t depth_lod, int y0, int y1);
int orc_gtao_filter(const OrcGtaoParams* p, const OrcImage* depth, const OrcImage* raw, const OrcImage* filtered,
                    uint32_t depth_lod, int y0, int y1);
int orc_gtao_accumulate(const OrcAccumParams* p, const OrcImage* depth, const OrcImage* prev_depth,
                        const OrcImage* filtered, const OrcImage* accumulated, const OrcImage* dvelocity,
                        const OrcImage* history, uint32_t depth_lod, int y0, int y1);

/* reference shaders/defered_shading/shader.frag, shaders/taa/resolve.comp */
int orc_deferred_shading(const OrcShadingParams* p, const OrcImage* albedo, const OrcImage* normal,
                         const OrcImage* material, const OrcImage* depth, const OrcImage* occlusion,
                         const OrcImage* brdf_lut, const OrcImage* reflections, const OrcImage* out_color, int y0, int y1);
int orc_taa_resolve(const OrcTaaParams* p, const OrcImage* history, const OrcImage* prev_depth, const OrcImage* depth,
                    const OrcImage* velocity, const OrcImage* color, const OrcImage* out_target, int y0, int y1);

/* secondary passes: reference shaders/ssao/shader.frag, shaders/ssr/shader.frag (+ include/screen_trace.glsl:51-100) */
int orc_ssao(const OrcSsaoParams* p, const OrcImage* depth, const OrcImage* out_r8, int y0, int y1);
int orc_ssr_legacy(const OrcSsrParams* p, const OrcImage* depth, const OrcImage* normal, const OrcImage* frame,
                   const OrcImage* material, const OrcImage* out_rgba8, int y0, int y1);

/* ScreenSpaceTrace: reference shaders/screen_trace/{trace,filter,accumulate}.comp */
int orc_screen_trace_main(const OrcScreenTraceParams* p, const OrcImage* depth, const OrcImage* normal, const OrcImage* color,
                          const OrcImage* material, const OrcImage* out_raw, int y0, int y1);
int orc_screen_trace_filter(float znear, float zfar, const OrcImage* raw, const OrcImage* depth, const OrcImage* out_filtered, int y0, int y1);
int orc_screen_trace_accumulate(const float* fovy_aspect_znear_zfar, const OrcImage* depth, const OrcImage* prev_depth,
                                const OrcImage* filtered, const OrcImage* accumulated, int y0, int y1);

/* GTAO variants: reference shaders/gtao/{main.frag,reproject.comp}, shaders/gtao_opt/{deinterleave,main_deinterleaved}.comp.
 * p->angle_offset is the push constant; the other OrcGtaoParams extras are ignored.  Array images: layer l at base + l * layer_stride. */
int orc_gtao_main_graphics(const OrcGtaoParams* p, const OrcImage* depth, const OrcImage* normal, const OrcImage* raw, uint32_t depth_lod, int y0, int y1);
int orc_gtao_deinterleave(const OrcImage* depth, uint32_t depth_lod, const OrcImage* out_r32f_array, uint64_t layer_stride, int pattern_step);
int orc_gtao_main_deinterleaved(const OrcGtaoParams* p, const OrcImage* depth_array, uint64_t layer_stride, const OrcImage* normal,
                                const OrcImage* raw, int pattern_n, uint32_t layer, int y0, int y1);
int orc_gtao_reproject(const float* fovy_aspect_znear_zfar, const OrcImage* depth, const OrcImage* prev_depth, uint32_t depth_lod,
                       const OrcImage* gtao, const OrcImage* prev_gtao, const OrcImage* out, int y0, int y1);

/* probe_renderer compute path: reference shaders/cube2oct/shader.comp, probe_downsample/shader.frag, trace_probe/shader.comp.
 * Cube maps: face f at base + f * face_stride; probe arrays: layer l at base + l * layer_stride. */
int orc_cube2oct(const OrcImage* cube_color, uint64_t color_face_stride, const OrcImage* cube_distance, uint64_t distance_face_stride,
                 const OrcImage* oct_color_layer, const OrcImage* oct_depth_layer);
int orc_probe_downsample(const OrcImage* probe_depth_layer);
int orc_trace_probe(const OrcProbeTraceParams* p, const OrcImage* depth, const OrcImage* normal, const OrcImage* probe_color,
                    uint64_t color_layer_stride, const OrcImage* probe_depth, uint64_t depth_layer_stride, const OrcImage* out, int y0, int y1);

/* SSSR tile path (dead in AdvancedSSR::run): reference shaders/advanced_ssr/{classification,trace_indirect,regression}.comp.
 * indirect = VkDispatchIndirectCommand {x, y, z}; classification appends to the lists in ascending tile index. */
int orc_sssr_classification(int tex_w, int tex_h, float g_max_roughness, float g_glossy_value, const OrcImage* material,
                            int32_t* reflective_tiles, int32_t* glossy_tiles, uint32_t* reflective_indirect, uint32_t* glossy_indirect);
int orc_sssr_trace_indirect(const OrcTraceParams* p, uint32_t reflection_type, const OrcImage* depth, const OrcImage* dnormal,
                            const OrcImage* material, const float* halton_vec4x128, const int32_t* tiles, uint32_t tile_count,
                            const OrcImage* out_rays);
int orc_sssr_tile_regression(const float* camera_to_world16, const float* fovy_aspect_znear_zfar, const OrcImage* depth,
                             const OrcImage* out_planes);

/* Capture path: reference src/util_passes.cpp:152-180 blit_image, src/image_readback.cpp:67-110, src/main.cpp:152-176. */
int orc_blit_image(const OrcImage* src, const OrcImage* dst);
int orc_readback_pack(const OrcImage* image, uint32_t mip, uint32_t mode, void* out_tight);

/* shared-math probes for the analytic anchors (tests only) */
void orc_probe_encode_normal(const float* v3, float* out2);
void orc_probe_decode_normal(const float* uv2, float* out3);
void orc_probe_reconstruct(const float* uv2, float d, const float* cam4, float* out3);
void orc_probe_project(const float* v3, const float* cam4, float* out3);
float orc_probe_half_roundtrip(float v);
float orc_probe_srgb_decode(uint32_t code);
float orc_probe_srgb_encode(float linear);

/* G-buffer producer conventions (passes_gbuf.cpp): the jitter table of reference src/main.cpp:93-108 and the vertex /
 * fragment arithmetic of shaders/gbuf/opaque_taa.vert:37-43, opaque_taa.frag:45 for one world-space point. */
void orc_taa_offset(uint32_t index, uint32_t w, uint32_t h, float* out_jitter4);
void orc_gbuf_point(const float* view_projection16, const float* prev_view_projection16, const float* jitter4,
                    const float* world_pos3, float* out_position4, float* out_velocity2);

#ifdef __cplusplus
}
#endif
#endif
