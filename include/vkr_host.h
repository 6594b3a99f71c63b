/* vkr_host.h — C API of the headless host (libvkr_host.so).
 *
 * The host is C++: a rendergraph shim plus the reference's pass classes (vk-renderer_b200/host/),
 * driven in the reference's frame order.  VkrApp is that frame loop without the window:
 *   create      = reference src/main.cpp:255-309  (RenderGraph, Gbuffer, DownsamplePass, GTAO(w,h,0,1),
 *                 AdvancedSSR, TAA, LUT pre-integration, DeferedShadingPass, color_out_tex,
 *                 clear_depth(prev_depth, 1.0))
 *   render_frame = reference src/main.cpp:329-421 (update_params, downsample_pass.run, ssr.run,
 *                 gtao.add_main/filter/accumulate, shading_pass.draw, taa_pass.run, submit, the five remaps)
 * The G-buffer rasterisation (scene_renderer.draw_taa) is the INPUT of the path: it is uploaded from
 * host memory (upload_gbuffer) or produced on the device by the synthetic ray-caster (synth_gbuffer).
 * This C API exists so Python (bench.py, tests) can drive the C++ host; a C++ user links the
 * classes in vk-renderer_b200/host/passes.hpp directly.
 * All calls return 0 on success, negative VkrStatus otherwise (vkr_host_last_error() has the text).
 */
#ifndef VKR_HOST_H_INCLUDED
#define VKR_HOST_H_INCLUDED
#include "vkr_passes.h"
#include "vkr_synth.h"

#ifdef __cplusplus
extern "C" {
#endif

typedef struct VkrApp VkrApp;

const char* vkr_host_last_error(void);

/* stream: a cudaStream_t to run on (e.g. the caller's), or NULL for a stream owned by the app. */
VkrApp* vkr_app_create(int device, uint32_t width, uint32_t height, void* stream);
void vkr_app_destroy(VkrApp* app);

/* Asynchronous host->device copy of one frame's G-buffer (albedo, normal, material, velocity, depth
 * mip 0), each laid out as vkr_image_layout() describes (mip 0 only: row_pitch[0] * height bytes).
 * Pinned host memory makes the copies truly asynchronous. */
int vkr_app_upload_gbuffer(VkrApp* app, const void* albedo, const void* normal, const void* material, const void* velocity, const void* depth);
/* Device-side producer for the same five images. */
int vkr_app_synth_gbuffer(VkrApp* app, const VkrSynthScene* scene, const VkrDrawTAAParams* params);

/* Bind caller-owned DEVICE images as the frame's G-buffer instead of copying (inputs already resident
 * in HBM).  Each must match the app's image (extent, format, mip count — depth carries its whole mip
 * chain, which DownsamplePass fills in place).  The app borrows them: after the depth<->prev_depth
 * remap the previous frame's depth image keeps serving as prev_depth, so keep it alive for one more frame. */
int vkr_app_bind_gbuffer(VkrApp* app, const VkrImage* albedo, const VkrImage* normal, const VkrImage* material, const VkrImage* velocity,
                         const VkrImage* depth);

/* Per-task device timing (cudaEvents around every task the rendergraph replays).  task_timing(1) resets
 * and enables; task_times() waits for the stream and returns the number of distinct tasks, their names
 * ('\n'-separated), accumulated milliseconds and launch counts. */
int vkr_app_task_timing(VkrApp* app, int enable);
int vkr_app_task_times(VkrApp* app, char* names_out, uint64_t names_capacity, double* ms_out, uint64_t* count_out, uint32_t max_tasks);

/* One frame of the chain on the current G-buffer.  gtao_angle_jitter replaces the reference's
 * rand()/RAND_MAX - 0.5 (src/gtao.cpp:111).  Asynchronous. */
int vkr_app_render_frame(VkrApp* app, const VkrDrawTAAParams* params, float gtao_angle_jitter);

/* Descriptor of a named image AS OF THE LAST RENDERED FRAME (the ping-pong remaps are undone in the
 * lookup): albedo normal material velocity depth prev_depth dn dv rays raw filtered accumulated_ao
 * reflections blurred color taa pdf brdf ssao ssr_legacy
 * st_raw st_filtered st_accumulated gv_raw gv_filtered gv_output gv_deinterleaved (layer 0) tile_planes readback_image. */
int vkr_app_get_image(VkrApp* app, const char* name, VkrImage* out);
/* Asynchronous device->host copy of the whole allocation of a named image (all mips). */
int vkr_app_readback(VkrApp* app, const char* name, void* host_dst, uint64_t capacity_bytes);
/* Asynchronous device->host copy of a named rendergraph buffer of the SSSR tile path (option alt.ssr_tiles; reference
 * src/advanced_ssr.cpp:77-83): reflective_indirect glossy_indirect (3 x uint32) reflective_tiles glossy_tiles (int32 each).
 * size_out (may be NULL) receives the buffer size; host_dst NULL only queries it. */
int vkr_app_readback_buffer(VkrApp* app, const char* name, void* host_dst, uint64_t capacity_bytes, uint64_t* size_out);
/* The capture buttons of the reference's frame loop (src/main.cpp:352-357,:393-412 + callbacks :118-176), synchronous:
 * kind 0 = "Final frame": blit_image(TAA output -> sRGB8 read-back image), ReadBackSystem::read_image, PNG with alpha 255;
 * kind 1 = depth mip 0 as a PNG of the 24-bit codes (R = low byte); kind 2 = the same codes as CSV ("y, 0,1,..", hex cells).
 * Tasks issued: ["CopyImage",] "ImageRead" (vkr_app_last_tasks). */
int vkr_app_capture(VkrApp* app, uint32_t kind, const char* path);
/* Host-only: 8-bit RGBA PNG from tightly packed rows (stands in for stbi_write_png(path, w, h, 4, data, 0)). */
int vkr_write_png_rgba8(const char* path, uint32_t width, uint32_t height, const void* rgba);
int vkr_app_sync(VkrApp* app);
void* vkr_app_stream(VkrApp* app);
/* Names of the tasks replayed by the last submit, '\n'-separated (for tests). */
const char* vkr_app_last_tasks(VkrApp* app);
/* One character per task of the last submit, in the same order: '0' = issued on the app's stream, '1' = on the rendergraph's
 * second stream (independent branches of a frame run side by side; option "graph.concurrent_branches", default 1; always '0'
 * while task timing is on). */
const char* vkr_app_last_lanes(VkrApp* app);
/* How the rendergraph's submits reached the GPU so far: out4 = {frames launched as ONE CUDA graph (captured per frame, applied
 * to the instantiated graph with cudaGraphExecUpdate), graph instantiations, submits issued eagerly (the first frame, frames
 * with a read-back task, task timing on, single-stream mode), captures that failed}.  Option "graph.frame_graph" (default 1)
 * switches the graph path.  Reference analogue: one command buffer per frame, src/rendergraph/rendergraph.cpp:271-337. */
int vkr_app_frame_graph_stats(VkrApp* app, uint64_t* out4);
/* The parameter blocks the optional "SSAO" / "SSR" tasks of the last frame were given (options alt.ssao / alt.ssr; the
 * passes of reference src/ssao.cpp and src/ssr.cpp, which the reference's frame loop does not schedule). */
int vkr_app_alt_params(VkrApp* app, VkrSsaoParams* ssao_out, VkrSsrParams* ssr_out);
/* ProbeRenderer::render_probe_grid (reference src/probe_renderer.cpp:253-283) on the app's 4x4 grid of 256^2 octahedral probes:
 * per probe the tasks CubemapProbe (synthetic faces, vkr_synth_cubemap), Cubemap2Octahedral, DownsampleProbe; submitted at once.
 * With option alt.trace_probe every later frame ends with ProbeTracePass::run into image "probe_reflection". */
int vkr_app_render_probe_grid(VkrApp* app, const VkrSynthScene* scene, const float* min3, const float* max3);
/* Descriptor of layer 0 of a named (array) image, the byte distance between its layers and their number:
 * probe_color probe_depth (16 layers) probe_cube_color probe_cube_distance (6 faces) and every name of vkr_app_get_image. */
int vkr_app_get_image_array(VkrApp* app, const char* name, VkrImage* out, uint64_t* layer_stride, uint32_t* layers);
/* Host-side state of the last frame by name: "gtao.angle_offset", "gv.angle_offset" (the angle push constants of the live GTAO
 * and of the variant object behind option alt.gtao_variants). */
int vkr_app_get_value(VkrApp* app, const char* name, float* out);
/* Likewise for the "ScreenTrace" task (option alt.screen_trace; reference src/screen_trace.cpp). */
int vkr_app_screen_trace_params(VkrApp* app, VkrScreenTraceParams* out);
/* alt.ssr_tiles: AdvancedSSR::run also issues SSSR_Clear, SSSR_Classification, SSSR_Tile_Regression and the indirect SSSR_trace
 * (reference src/advanced_ssr.cpp:547-550, commented out there); ssr.glossy_roughness_value = the classification threshold. */
/* Runtime switches mirroring the ImGui state of the reference passes. fused_hiz: one launch for the
 * whole DownsamplePass instead of the per-task calls.  graph.concurrent_branches: see vkr_app_last_lanes.
 * graph.frame_graph: see vkr_app_frame_graph_stats. */
int vkr_app_set_option(VkrApp* app, const char* name, float value);

#ifdef __cplusplus
}
#endif
#endif
