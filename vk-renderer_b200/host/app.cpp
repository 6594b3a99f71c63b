// app.cpp — the reference's frame loop without the window (see include/vkr_host.h), and its C API.
#include <cuda_runtime.h>

#include <cmath>
#include <cstdarg>
#include <cstdio>
#include <map>
#include <string>

#include "../../include/vkr_host.h"
#include "passes.hpp"
#include "readback.hpp"

using namespace rendergraph;

static thread_local char g_host_err[512] = "";
static void host_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_host_err, sizeof(g_host_err), fmt, ap);
  va_end(ap);
}

static void copy16(float* dst, const float* src) { std::memcpy(dst, src, 64); }

// glm::perspectiveRH_ZO in fp32 (the projection Camera::get_projection hands to SSAOInParams)
static glm::mat4 perspective(float fovy, float aspect, float zn, float zf) {
  const float t = std::tan(fovy / 2.f);
  glm::mat4 p;
  std::memset(p.m, 0, sizeof(p.m));
  p.m[0] = 1.f / (aspect * t); p.m[5] = 1.f / t; p.m[10] = zf / (zn - zf); p.m[11] = -1.f; p.m[14] = -(zf * zn) / (zf - zn);
  return p;
}

struct VkrApp {
  uint32_t width, height;
  RenderGraph render_graph;
  Gbuffer gbuffer;
  DownsamplePass downsample_pass;
  GTAO gtao;
  AdvancedSSR ssr;
  TAA taa_pass;
  DeferedShadingPass shading_pass;
  ImageResourceId shadows_tex, color_out_tex;
  // the earlier AO / reflection passes (src/ssao.cpp, src/ssr.cpp): in the reference tree but not in its frame loop, so off by default
  ImageResourceId ssao_tex, ssr_tex;
  SSAOPass ssao_pass;
  ScreenSpaceTrace screen_trace;
  ProbeRenderer probe_renderer;                      // src/probe_renderer.hpp: not instantiated by the reference's main(), off by default
  OctahedralProbeGrid probe_grid;
  ProbeTracePass probe_trace;
  ImageResourceId probe_reflection;
  // capture path (main.cpp:220-223,:247,:352-357,:393-396): sRGB8 read-back image the TAA output is blitted into + ReadBackSystem
  ImageResourceId readback_image;
  ReadBackSystem readback_system;
  bool alt_trace_probe = false, probes_rendered = false;
  GTAO gtao_alt;     // a second GTAO object so the variants do not disturb the live chain's images
  bool alt_ssao = false, alt_ssr = false, alt_screen_trace = false, alt_gtao_variants = false;
  VkrSsrParams last_ssr{};
  std::string task_names, task_lanes;
  uint64_t frames = 0;

  VkrApp(int device, uint32_t w, uint32_t h, void* stream)
      : width{w}, height{h}, render_graph{device, stream}, gbuffer{render_graph, w, h}, downsample_pass{},
        gtao{render_graph, w, h, /*USE_RAY_QUERY*/ false, /*half_res*/ true},                   // main.cpp:265
        ssr{render_graph, w, h}, taa_pass{render_graph, w, h}, shading_pass{render_graph},
        ssao_tex{create_ssao_texture(render_graph, w, h)}, ssr_tex{create_ssr_tex(render_graph, w, h)}, ssao_pass{render_graph, ssao_tex, 0},
        screen_trace{render_graph, w, h}, probe_renderer{render_graph, CUBE_SIZE}, probe_grid{render_graph, 4, PROBE_SIZE},
        gtao_alt{render_graph, w, h, false, true, 2} {
    { gpu::ImageInfo pr; pr.format = VKR_FORMAT_R8G8B8A8_UNORM; pr.width = w; pr.height = h; probe_reflection = render_graph.create_image(pr); }
    { gpu::ImageInfo ri; ri.format = VKR_FORMAT_R8G8B8A8_SRGB; ri.width = w; ri.height = h; readback_image = render_graph.create_image(ri); }   // main.cpp:220-223,:278
    ssr.preintegrate_pdf(render_graph);                                                           // main.cpp:269-270
    ssr.preintegrate_brdf(render_graph);
    gpu::ImageInfo sh; sh.format = VKR_FORMAT_D24_UNORM_S8_UINT; sh.width = 16; sh.height = 16;  // bound, never sampled (main.cpp:276-284)
    shadows_tex = render_graph.create_image(sh);
    gpu::ImageInfo co; co.format = VKR_FORMAT_R8G8B8A8_SRGB; co.width = w; co.height = h;        // main.cpp:289-291
    color_out_tex = render_graph.create_image(co);
    render_graph.submit();                                                                        // main.cpp:304 — LUTs are built here
    // clear_depth(render_graph, gbuffer.prev_depth) -> 1.0 on every mip (main.cpp:306, util_passes.cpp:93)
    // memset 0xFF also sets the stencil byte, which every kernel masks off (depth = low 24 bits).
    const VkrImage& pd = render_graph.image(gbuffer.prev_depth);
    VkrImage tmp;
    const uint64_t total = vkr_image_layout(pd.width, pd.height, pd.format, pd.mip_count, &tmp);
    if (cudaMemsetAsync(pd.base, 0xFF, total, (cudaStream_t)render_graph.stream()) != cudaSuccess)
      throw std::runtime_error("clear_depth(prev_depth) failed");
  }

  void render_frame(const DrawTAAParams& draw_params, float angle_jitter) {
    const float fovy = draw_params.fovy_aspect_znear_zfar.x, aspect = draw_params.fovy_aspect_znear_zfar.y;
    const float znear = draw_params.fovy_aspect_znear_zfar.z, zfar = draw_params.fovy_aspect_znear_zfar.w;
    glm::mat4 shadow_mvp;                                                                         // unused by the shader
    shading_pass.update_params(draw_params.camera, shadow_mvp, fovy, aspect, znear, zfar);        // main.cpp:339

    downsample_pass.run(render_graph, gbuffer.normal, gbuffer.velocity_vectors, gbuffer.depth, gbuffer.downsampled_normals,
                        gbuffer.downsampled_velocity_vectors);                                     // main.cpp:347

    auto normal_mat = glm::transpose(glm::inverse(draw_params.camera));                           // main.cpp:368
    GTAOParams gtao_params{normal_mat, fovy, aspect, znear, zfar};
    AdvancedSSRParams assr_params{normal_mat, fovy, aspect, znear, zfar};

    ssr.run(render_graph, assr_params, draw_params, gbuffer, gtao.raw);                           // main.cpp:375
    gtao.angle_jitter = angle_jitter;
    gtao.add_main_pass(render_graph, gtao_params, gbuffer.depth, gbuffer.normal, gbuffer.material, ssr.get_preintegrated_pdf());  // :384
    gtao.add_filter_pass(render_graph, gtao_params, gbuffer.depth);                               // :387
    gtao.add_accumulate_pass(render_graph, draw_params, gbuffer);                                 // :388
    shading_pass.draw(render_graph, gbuffer, shadows_tex, gtao.accumulated_ao, ssr.get_preintegrated_brdf(), ssr.get_blurred(), color_out_tex);  // :390
    taa_pass.run(render_graph, gbuffer, color_out_tex, draw_params);                              // :391
    if (alt_ssao) {
      SSAOInParams sp{perspective(fovy, aspect, znear, zfar), fovy, aspect, znear, zfar};
      ssao_pass.draw(render_graph, gbuffer.depth, ssao_tex, sp);
    }
    if (alt_ssr) {
      SSRParams rp{normal_mat, fovy, aspect, znear, zfar};
      copy16(last_ssr.normal_mat, normal_mat.m);
      last_ssr.fovy = fovy; last_ssr.aspect = aspect; last_ssr.znear = znear; last_ssr.zfar = zfar;
      add_ssr_pass(render_graph, gbuffer.depth, gbuffer.normal, color_out_tex, gbuffer.material, ssr_tex, rp);
    }
    if (alt_screen_trace) {
      ScreenTraceParams tp{normal_mat, fovy, aspect, znear, zfar};
      screen_trace.next_angle_random = angle_jitter + 0.5f;        // the same uniform the GTAO angle uses this frame
      screen_trace.next_random_offset = 0.25f * (float)(frames & 3);
      screen_trace.add_main_pass(render_graph, tp, gbuffer.depth, gbuffer.normal, color_out_tex, gbuffer.material);
      screen_trace.add_filter_pass(render_graph, tp, gbuffer.depth);
      screen_trace.add_accumulate_pass(render_graph, tp, gbuffer.depth, gbuffer.prev_depth);
    }
    if (alt_trace_probe) {
      if (!probes_rendered) throw std::runtime_error{"alt.trace_probe: call vkr_app_render_probe_grid first"};
      ProbeTraceParams pp{glm::inverse(draw_params.camera), fovy, aspect, znear, zfar};
      probe_trace.run(render_graph, probe_grid, gbuffer.depth, gbuffer.normal, probe_reflection, pp);
    }
    if (alt_gtao_variants) {
      gtao_alt.angle_jitter = angle_jitter;
      gtao_alt.add_main_pass_graphics(render_graph, gtao_params, gbuffer.depth, gbuffer.normal);
      gtao_alt.add_filter_pass(render_graph, gtao_params, gbuffer.depth);
      GTAOReprojection rp{glm::mat4{}, fovy, aspect, znear, zfar};
      gtao_alt.add_reprojection_pass(render_graph, rp, gbuffer.depth, gbuffer.prev_depth);
      gtao_alt.deinterleave_depth(render_graph, gbuffer.depth);
      gtao_alt.add_main_pass_deinterleaved(render_graph, gtao_params, gbuffer.normal);
    }
    render_graph.submit();                                                                        // :402
    task_names.clear();
    for (auto& n : render_graph.last_task_names()) { task_names += n; task_names += '\n'; }
    task_lanes.clear();
    for (int l : render_graph.last_task_lanes()) task_lanes += (char)('0' + l);

    render_graph.remap(gbuffer.depth, gbuffer.prev_depth);                                        // main.cpp:416-420
    render_graph.remap(gtao.output, gtao.prev_frame);
    if (alt_gtao_variants) render_graph.remap(gtao_alt.output, gtao_alt.prev_frame);
    taa_pass.remap_targets(render_graph);
    ssr.remap_images(render_graph);
    gtao.remap(render_graph);
    frames++;
  }

  // Image by name as of the last rendered frame: after the remaps the frame's depth is in the
  // prev_depth slot, its TAA target in the history slot, etc.
  bool lookup(const std::string& name, ImageResourceId& id) {
    const bool r = frames > 0;   // remapped at least once
    if (name == "albedo") id = gbuffer.albedo;
    else if (name == "normal") id = gbuffer.normal;
    else if (name == "material") id = gbuffer.material;
    else if (name == "velocity") id = gbuffer.velocity_vectors;
    else if (name == "depth") id = r ? gbuffer.prev_depth : gbuffer.depth;
    else if (name == "prev_depth") id = r ? gbuffer.depth : gbuffer.prev_depth;
    else if (name == "dn") id = gbuffer.downsampled_normals;
    else if (name == "dv") id = gbuffer.downsampled_velocity_vectors;
    else if (name == "rays") id = ssr.get_rays();
    else if (name == "raw") id = gtao.raw;
    else if (name == "filtered") id = gtao.filtered;
    else if (name == "accumulated_ao") id = r ? gtao.accumulated_history : gtao.accumulated_ao;
    else if (name == "reflections") id = ssr.get_ouput();
    else if (name == "blurred") id = r ? ssr.get_blurred_history() : ssr.get_blurred();
    else if (name == "color") id = color_out_tex;
    else if (name == "taa") id = r ? taa_pass.get_history() : taa_pass.get_output();
    else if (name == "pdf") id = ssr.get_preintegrated_pdf();
    else if (name == "brdf") id = ssr.get_preintegrated_brdf();
    else if (name == "ssao") id = ssao_tex;
    else if (name == "ssr_legacy") id = ssr_tex;
    else if (name == "probe_color") id = probe_grid.color_array;
    else if (name == "probe_depth") id = probe_grid.depth_array;
    else if (name == "probe_cube_color") id = probe_renderer.get_cubemap_color();
    else if (name == "probe_cube_distance") id = probe_renderer.get_cubemap_distance();
    else if (name == "probe_reflection") id = probe_reflection;
    else if (name == "gv_raw") id = gtao_alt.raw;
    else if (name == "gv_filtered") id = gtao_alt.filtered;
    else if (name == "gv_output") id = r ? gtao_alt.prev_frame : gtao_alt.output;
    else if (name == "gv_deinterleaved") id = gtao_alt.deinterleaved_depth;
    else if (name == "st_raw") id = screen_trace.raw;
    else if (name == "st_filtered") id = screen_trace.filtered;
    else if (name == "st_accumulated") id = screen_trace.accumulated;
    else if (name == "tile_planes") id = ssr.get_tile_planes();
    else if (name == "readback_image") id = readback_image;
    else return false;
    return true;
  }
};

extern "C" {

const char* vkr_host_last_error(void) { return g_host_err; }

#define VKR_TRY(...)                                                         \
  try { __VA_ARGS__; return VKR_OK; }                                        \
  catch (const std::exception& e) { host_error("%s", e.what()); return VKR_ERR_CUDA; }

VkrApp* vkr_app_create(int device, uint32_t width, uint32_t height, void* stream) {
  try {
    if (width < 16 || height < 16) { host_error("vkr_app_create: extent too small"); return nullptr; }
    return new VkrApp(device, width, height, stream);
  } catch (const std::exception& e) {
    host_error("%s", e.what());
    return nullptr;
  }
}

void vkr_app_destroy(VkrApp* app) { delete app; }

int vkr_app_upload_gbuffer(VkrApp* app, const void* albedo, const void* normal, const void* material, const void* velocity, const void* depth) {
  if (!app || !albedo || !normal || !material || !velocity || !depth) { host_error("upload_gbuffer: null argument"); return VKR_ERR_INVALID_ARGUMENT; }
  VKR_TRY({
    cudaStream_t st = (cudaStream_t)app->render_graph.stream();
    const void* src[5] = {albedo, normal, material, velocity, depth};
    ImageResourceId ids[5] = {app->gbuffer.albedo, app->gbuffer.normal, app->gbuffer.material, app->gbuffer.velocity_vectors, app->gbuffer.depth};
    for (int i = 0; i < 5; i++) {
      const VkrImage& im = app->render_graph.image(ids[i]);
      size_t bytes = (size_t)im.row_pitch[0] * im.height;
      if (cudaMemcpyAsync(im.base, src[i], bytes, cudaMemcpyHostToDevice, st) != cudaSuccess) throw std::runtime_error("upload_gbuffer: cudaMemcpyAsync failed");
    }
  })
}

int vkr_app_synth_gbuffer(VkrApp* app, const VkrSynthScene* scene, const VkrDrawTAAParams* params) {
  if (!app || !scene || !params) { host_error("synth_gbuffer: null argument"); return VKR_ERR_INVALID_ARGUMENT; }
  VKR_TRY({
    auto& g = app->render_graph;
    int rc = vkr_synth_gbuffer(scene, params, &g.image(app->gbuffer.albedo), &g.image(app->gbuffer.normal), &g.image(app->gbuffer.material),
                               &g.image(app->gbuffer.velocity_vectors), &g.image(app->gbuffer.depth), g.stream());
    vkr_check(rc, "GbufferPass(synthetic)");
  })
}

int vkr_app_render_frame(VkrApp* app, const VkrDrawTAAParams* p, float gtao_angle_jitter) {
  if (!app || !p) { host_error("render_frame: null argument"); return VKR_ERR_INVALID_ARGUMENT; }
  VKR_TRY({
    DrawTAAParams d;
    d.mvp = glm::mat4(p->mvp); d.prev_mvp = glm::mat4(p->prev_mvp); d.camera = glm::mat4(p->camera); d.prev_camera = glm::mat4(p->prev_camera);
    d.jitter = {p->jitter[0], p->jitter[1], p->jitter[2], p->jitter[3]};
    d.fovy_aspect_znear_zfar = {p->fovy_aspect_znear_zfar[0], p->fovy_aspect_znear_zfar[1], p->fovy_aspect_znear_zfar[2], p->fovy_aspect_znear_zfar[3]};
    app->render_frame(d, gtao_angle_jitter);
  })
}

int vkr_app_bind_gbuffer(VkrApp* app, const VkrImage* albedo, const VkrImage* normal, const VkrImage* material, const VkrImage* velocity, const VkrImage* depth) {
  if (!app || !albedo || !normal || !material || !velocity || !depth) { host_error("bind_gbuffer: null argument"); return VKR_ERR_INVALID_ARGUMENT; }
  VKR_TRY({
    auto& g = app->render_graph;
    g.rebind_image(app->gbuffer.albedo, *albedo);
    g.rebind_image(app->gbuffer.normal, *normal);
    g.rebind_image(app->gbuffer.material, *material);
    g.rebind_image(app->gbuffer.velocity_vectors, *velocity);
    g.rebind_image(app->gbuffer.depth, *depth);
  })
}

int vkr_app_task_timing(VkrApp* app, int enable) {
  if (!app) return VKR_ERR_INVALID_ARGUMENT;
  VKR_TRY({ app->render_graph.reset_task_times(); app->render_graph.enable_task_timing(enable != 0); })
}

int vkr_app_task_times(VkrApp* app, char* names_out, uint64_t names_capacity, double* ms_out, uint64_t* count_out, uint32_t max_tasks) {
  if (!app || !names_out || !ms_out || !count_out) return VKR_ERR_INVALID_ARGUMENT;
  try {
    const auto& t = app->render_graph.collect_task_times();
    std::string names;
    uint32_t n = 0;
    for (; n < t.size() && n < max_tasks; n++) { names += t[n].name; names += '\n'; ms_out[n] = t[n].ms; count_out[n] = t[n].count; }
    if (names.size() + 1 > names_capacity) { host_error("task_times: names buffer too small"); return VKR_ERR_INVALID_ARGUMENT; }
    std::memcpy(names_out, names.c_str(), names.size() + 1);
    return (int)n;
  } catch (const std::exception& e) { host_error("%s", e.what()); return VKR_ERR_CUDA; }
}

int vkr_app_get_image(VkrApp* app, const char* name, VkrImage* out) {
  if (!app || !name || !out) { host_error("get_image: null argument"); return VKR_ERR_INVALID_ARGUMENT; }
  ImageResourceId id;
  if (!app->lookup(name, id)) { host_error("get_image: unknown image '%s'", name); return VKR_ERR_INVALID_ARGUMENT; }
  *out = app->render_graph.image(id);
  return VKR_OK;
}

int vkr_app_readback(VkrApp* app, const char* name, void* host_dst, uint64_t capacity) {
  VkrImage im;
  int rc = vkr_app_get_image(app, name, &im);
  if (rc) return rc;
  VkrImage tmp;
  uint64_t total = vkr_image_layout(im.width, im.height, im.format, im.mip_count, &tmp);
  if (!host_dst || capacity < total) { host_error("readback: destination too small (%llu < %llu)", (unsigned long long)capacity, (unsigned long long)total); return VKR_ERR_INVALID_ARGUMENT; }
  if (cudaMemcpyAsync(host_dst, im.base, total, cudaMemcpyDeviceToHost, (cudaStream_t)app->render_graph.stream()) != cudaSuccess) {
    host_error("readback: cudaMemcpyAsync failed");
    return VKR_ERR_CUDA;
  }
  return VKR_OK;
}

int vkr_app_readback_buffer(VkrApp* app, const char* name, void* host_dst, uint64_t capacity, uint64_t* size_out) {
  if (!app || !name) { host_error("readback_buffer: null argument"); return VKR_ERR_INVALID_ARGUMENT; }
  const std::string n{name};
  rendergraph::BufferResourceId id;
  if (n == "reflective_indirect") id = app->ssr.get_reflective_indirect();
  else if (n == "glossy_indirect") id = app->ssr.get_glossy_indirect();
  else if (n == "reflective_tiles") id = app->ssr.get_reflective_tiles();
  else if (n == "glossy_tiles") id = app->ssr.get_glossy_tiles();
  else { host_error("readback_buffer: unknown buffer '%s'", name); return VKR_ERR_INVALID_ARGUMENT; }
  const uint64_t total = app->render_graph.buffer_size(id);
  if (size_out) *size_out = total;
  if (!host_dst) return VKR_OK;
  if (capacity < total) { host_error("readback_buffer: destination too small (%llu < %llu)", (unsigned long long)capacity, (unsigned long long)total); return VKR_ERR_INVALID_ARGUMENT; }
  if (cudaMemcpyAsync(host_dst, app->render_graph.buffer(id), total, cudaMemcpyDeviceToHost, (cudaStream_t)app->render_graph.stream()) != cudaSuccess) {
    host_error("readback_buffer: cudaMemcpyAsync failed");
    return VKR_ERR_CUDA;
  }
  return VKR_OK;
}

int vkr_app_capture(VkrApp* app, uint32_t kind, const char* path) {
  if (!app || !path || kind > 2) { host_error("capture: null argument or unknown kind"); return VKR_ERR_INVALID_ARGUMENT; }
  VKR_TRY({
    ImageResourceId src;
    ReadBackID id;
    if (kind == 0) {                                                                              // "Final frame", main.cpp:393-396
      app->lookup("taa", src);
      blit_image(app->render_graph, src, app->readback_image);
      id = app->readback_system.read_image(app->render_graph, app->readback_image, 0, 0, VKR_READBACK_OPAQUE_ALPHA);
    } else {                                                                                      // depth, main.cpp:354
      app->lookup("depth", src);
      id = app->readback_system.read_image(app->render_graph, src, 0, 0, VKR_READBACK_DEPTH24);
    }
    app->render_graph.submit();
    app->task_names.clear();
    for (auto& n : app->render_graph.last_task_names()) { app->task_names += n; app->task_names += '\n'; }
    uint32_t polls = 0;
    while (!app->readback_system.is_data_available(id)) {                                         // main.cpp:404-412, once per frame there
      app->readback_system.after_submit(app->render_graph);
      if (++polls > 16) throw std::runtime_error{"capture: read-back never became available"};
    }
    auto data = app->readback_system.get_data(id);
    const bool ok = kind == 0 ? get_rgba_cb(std::move(data), path) : kind == 1 ? get_depth_png_cb(std::move(data), path) : get_depth_cb(std::move(data), path);
    if (!ok) throw std::runtime_error{std::string{"capture: cannot write "} + path};
  })
}

int vkr_write_png_rgba8(const char* path, uint32_t width, uint32_t height, const void* rgba) {
  if (!write_png_rgba8(path, width, height, (const uint8_t*)rgba)) { host_error("write_png_rgba8: cannot write '%s'", path ? path : "(null)"); return VKR_ERR_INVALID_ARGUMENT; }
  return VKR_OK;
}

int vkr_app_sync(VkrApp* app) {
  if (!app) return VKR_ERR_INVALID_ARGUMENT;
  VKR_TRY(app->render_graph.wait_idle())
}

int vkr_app_render_probe_grid(VkrApp* app, const VkrSynthScene* scene, const float* min3, const float* max3) {
  if (!app || !scene || !min3 || !max3) { host_error("render_probe_grid: null argument"); return VKR_ERR_INVALID_ARGUMENT; }
  VKR_TRY({
    app->probe_renderer.render_probe_grid(app->render_graph, *scene, vec3{min3[0], min3[1], min3[2]}, vec3{max3[0], max3[1], max3[2]}, app->probe_grid);
    app->render_graph.submit();
    app->task_names.clear();
    for (auto& n : app->render_graph.last_task_names()) { app->task_names += n; app->task_names += '\n'; }
    app->task_lanes.clear();
    for (int l : app->render_graph.last_task_lanes()) app->task_lanes += (char)('0' + l);
    app->probes_rendered = true;
  })
}

int vkr_app_get_image_array(VkrApp* app, const char* name, VkrImage* out, uint64_t* layer_stride, uint32_t* layers) {
  if (!app || !name || !out || !layer_stride || !layers) { host_error("get_image_array: null argument"); return VKR_ERR_INVALID_ARGUMENT; }
  ImageResourceId id;
  if (!app->lookup(name, id)) { host_error("get_image_array: unknown image '%s'", name); return VKR_ERR_INVALID_ARGUMENT; }
  *out = app->render_graph.image(id);
  *layer_stride = app->render_graph.layer_stride(id);
  *layers = app->render_graph.get_descriptor(id).array_layers;
  return VKR_OK;
}

int vkr_app_get_value(VkrApp* app, const char* name, float* out) {
  if (!app || !name || !out) { host_error("get_value: null argument"); return VKR_ERR_INVALID_ARGUMENT; }
  std::string n = name;
  if (n == "gtao.angle_offset") *out = app->gtao.last_angle_offset;
  else if (n == "gv.angle_offset") *out = app->gtao_alt.last_angle_offset;
  else { host_error("get_value: unknown value '%s'", name); return VKR_ERR_INVALID_ARGUMENT; }
  return VKR_OK;
}

int vkr_app_screen_trace_params(VkrApp* app, VkrScreenTraceParams* out) {
  if (!app || !out) { host_error("screen_trace_params: null argument"); return VKR_ERR_INVALID_ARGUMENT; }
  *out = app->screen_trace.last_block;
  return VKR_OK;
}

int vkr_app_alt_params(VkrApp* app, VkrSsaoParams* ssao_out, VkrSsrParams* ssr_out) {
  if (!app) { host_error("alt_params: null app"); return VKR_ERR_INVALID_ARGUMENT; }
  if (ssao_out) *ssao_out = app->ssao_pass.last_block;
  if (ssr_out) *ssr_out = app->last_ssr;
  return VKR_OK;
}

void* vkr_app_stream(VkrApp* app) { return app ? app->render_graph.stream() : nullptr; }
const char* vkr_app_last_tasks(VkrApp* app) { return app ? app->task_names.c_str() : ""; }
const char* vkr_app_last_lanes(VkrApp* app) { return app ? app->task_lanes.c_str() : ""; }

int vkr_app_frame_graph_stats(VkrApp* app, uint64_t* out4) {
  if (!app || !out4) return VKR_ERR_INVALID_ARGUMENT;
  const auto& st = app->render_graph.frame_graph_stats();
  out4[0] = st.launches; out4[1] = st.instantiations; out4[2] = st.eager_submits; out4[3] = st.capture_failures;
  return VKR_OK;
}

int vkr_app_set_option(VkrApp* app, const char* name, float v) {
  if (!app || !name) return VKR_ERR_INVALID_ARGUMENT;
  std::string n = name;
  if (n == "fused_hiz") app->downsample_pass.fused = v != 0.f;
  else if (n == "graph.concurrent_branches") app->render_graph.enable_concurrent_branches(v != 0.f);
  else if (n == "graph.frame_graph") app->render_graph.enable_frame_graph(v != 0.f);
  else if (n == "gtao.mis") app->gtao.mis_gtao = v != 0.f;
  else if (n == "gtao.two_directions") app->gtao.two_directions = v != 0.f;
  else if (n == "gtao.only_reflections") app->gtao.only_reflections = v != 0.f;
  else if (n == "gtao.clear_history") app->gtao.clear_history = v != 0.f;
  else if (n == "gtao.weight_ratio") app->gtao.weight_ratio = v;
  else if (n == "ssr.max_roughness") app->ssr.settings.max_rougness = v;
  else if (n == "ssr.normalize") app->ssr.settings.normalize_reflections = v != 0.f;
  else if (n == "ssr.accumulate") app->ssr.settings.accumulate_reflections = v != 0.f;
  else if (n == "ssr.bilateral") app->ssr.settings.bilateral_filter = v != 0.f;
  else if (n == "ssr.use_blur") app->ssr.settings.use_blur = v != 0.f;
  else if (n == "ssr.max_accumulated_rays") app->ssr.settings.max_accumulated_rays = (int)v;
  else if (n == "alt.ssao") app->alt_ssao = v != 0.f;
  else if (n == "alt.ssr") app->alt_ssr = v != 0.f;
  else if (n == "alt.screen_trace") app->alt_screen_trace = v != 0.f;
  else if (n == "alt.gtao_variants") app->alt_gtao_variants = v != 0.f;
  else if (n == "alt.trace_probe") app->alt_trace_probe = v != 0.f;
  else if (n == "alt.ssr_tiles") app->ssr.settings.run_tile_path = v != 0.f;
  else if (n == "ssr.glossy_roughness_value") app->ssr.settings.glossy_roughness_value = v;
  else if (n == "alt.ssao_reference_packing") app->ssao_pass.reference_packing = v != 0.f;
  else if (n == "shading.only_ao") app->shading_pass.only_ao = v != 0.f;
  else if (n == "shading.min_roughness") app->shading_pass.min_max_roughness.x = v;
  else if (n == "shading.max_roughness") app->shading_pass.min_max_roughness.y = v;
  else { host_error("set_option: unknown option '%s'", name); return VKR_ERR_INVALID_ARGUMENT; }
  return VKR_OK;
}

}  // extern "C"
