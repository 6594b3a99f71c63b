// ORACLE — TEST INFRASTRUCTURE ONLY (see oracle.h header).  PARITY UNPINNED.
//
// The GTAO variants that sit beside the live compute path:
//   add_main_pass_graphics       reference src/gtao.cpp:349-413, shaders/gtao/main.frag:47-50,:69-93,:167-196
//   deinterleave_depth           reference src/gtao.cpp:445-472, shaders/gtao_opt/deinterleave.comp:10-21
//   add_main_pass_deinterleaved  reference src/gtao.cpp:474-526, shaders/gtao_opt/main_deinterleaved.comp:37-124
//   add_reprojection_pass        reference src/gtao.cpp:241-284, shaders/gtao/reproject.comp:26-66 (STATIC_REPROJECT, :7)
//
// As written in the reference (restated, not repaired):
//   * deinterleave dispatches over the OUTPUT extent (gtao.cpp:468), so only the top-left (w/2^n x h/2^n) texels of the depth
//     level are scattered, each layer receiving a (w/4^n x h/4^n) corner;
//   * main_deinterleaved loops over the array layers of its OUTPUT image `raw` (one layer, gtao.cpp:520) and dispatches over
//     raw's extent, so it evaluates layer 0 only, at uv = invocation / covered extent, and stores to pixel 2^n * invocation:
//     stores beyond raw's extent are dropped.  `layer` is an input here; the host class loops like the reference.
// The slice angle is evaluated in double and rounded once (both implementations march from the same direction bits).
#include <cmath>
#include "oracle.h"
#include "shader_include.hpp"

using namespace orc;

namespace {

const float MAX_THIKNESS = 0.1f;                                                                  // main.frag:67
const int SAMPLES = 20;                                                                           // :32
float sin_r(float x) { return (float)std::sin((double)x); }
float cos_r(float x) { return (float)std::cos((double)x); }
float gtao_direction(ivec2 pos) { return (1.0f / 16.0f) * (float)((((pos.x + pos.y) & 3) << 2) + (pos.x & 3)); }

// main.frag:69-93 / main_deinterleaved.comp:54-77 (the depth source is the only difference)
float find_horizon(const View& depth, vec2 start, vec3 camera_start, vec2 dir, int samples_count, vec3 v, const Cam& cam) {
  float h_cos = -1.0f;
  float previous_z = camera_start.z;
  for (int i = 1; i <= samples_count; i++) {
    vec2 tc = start + ((float)i / samples_count) * dir;
    float sample_depth = textureLod(depth, tc, 0).x;
    vec3 sample_pos = reconstruct_view_vec(tc, sample_depth, cam);
    if (sample_pos.z > previous_z + MAX_THIKNESS) break;
    previous_z = sample_pos.z;
    vec3 sample_offset = sample_pos - camera_start;
    float sample_cos = dot(v, normalize(sample_offset));
    if (sample_cos > h_cos) h_cos = sample_cos;
  }
  return h_cos;
}

// main.frag:167-196 / main_deinterleaved.comp:87-116, dirs_count = 1
float gtao_camera_space(const View& depth, const View& gbuffer_normal, const mat4& normal_mat, const Cam& cam, float angle_offset, ivec2 pos, vec2 screen_uv) {
  float frag_depth = texture(depth, screen_uv).x;
  if (frag_depth >= 1.f) return 1.f;
  vec3 camera_pos = reconstruct_view_vec(screen_uv, frag_depth, cam);
  vec3 w0 = -normalize(camera_pos);
  vec3 camera_normal = normalize((normal_mat * vec4(decode_normal(texture(gbuffer_normal, screen_uv).xy()), 0)).xyz());
  vec2 dir_radius = min(200.0f / length(camera_pos), 32.0f) / to_vec2(textureSize(depth, 0));
  float base_angle = gtao_direction(pos) + angle_offset;
  float sum = 0.f;
  {
    float angle = 2 * PI * (base_angle + 0.0f / 1.0f);
    vec2 sample_direction = dir_radius * vec2(cos_r(angle), sin_r(angle));
    vec3 sample_end_pos = reconstruct_view_vec(screen_uv + sample_direction, frag_depth, cam);
    vec3 slice_normal = normalize(cross(w0, -sample_end_pos));
    vec3 normal_projected = camera_normal - dot(camera_normal, slice_normal) * slice_normal;
    float n = PI / 2.0f - std::acos(dot(normalize(normal_projected), normalize(sample_end_pos - camera_pos)));
    float h_cos = find_horizon(depth, screen_uv, camera_pos, sample_direction, SAMPLES, w0, cam);
    float h = std::acos(h_cos);
    h = min(n + min(h - n, PI / 2.0f), h);
    sum += length(normal_projected) * 0.25f * max(-std::cos(2 * h - n) + std::cos(n) + 2 * h * std::sin(n), 0.0f);
  }
  return 2 * sum / 1.0f;
}

// One array layer of `im` as an image of its own.
Image layer_image(const Image& im, uint64_t layer_stride, uint32_t layer) {
  Image l = im;
  l.base = (uint8_t*)im.base + layer_stride * layer;
  return l;
}

}  // namespace

// Fragment pass over a fullscreen triangle: every pixel of raw, uv at pixel centres.  The float output lands in raw.r; the
// other channels of the RGBA16F attachment are undefined in the reference and written as (0, 0, 1) here.
extern "C" int orc_gtao_main_graphics(const OrcGtaoParams* p, const OrcImage* depth_, const OrcImage* normal_, const OrcImage* raw_,
                                      uint32_t depth_lod, int y0, int y1) {
  const Image& raw = *(const Image*)raw_;
  View depth(*(const Image*)depth_, (int)depth_lod, 1), normal(*(const Image*)normal_);
  const mat4 normal_mat(p->normal_mat);
  const Cam cam(p->fovy, p->aspect, p->znear, p->zfar);
  const int W = raw.width, H = raw.height;
  if (y0 < 0) y0 = 0;
  if (y1 <= 0 || y1 > H) y1 = H;
#pragma omp parallel for schedule(dynamic, 4)
  for (int py = y0; py < y1; py++)
    for (int px = 0; px < W; px++) {
      vec2 screen_uv = (vec2((float)px, (float)py) + vec2(0.5f, 0.5f)) / vec2((float)W, (float)H);
      float occlusion = gtao_camera_space(depth, normal, normal_mat, cam, p->angle_offset, ivec2(px, py), screen_uv);
      imageStore(raw, 0, ivec2(px, py), vec4(occlusion, 0, 0, 1));
    }
  return 0;
}

// out: R32_SFLOAT array image, layer l at base + l * layer_stride, (1 << pattern_n)^2 layers.
extern "C" int orc_gtao_deinterleave(const OrcImage* depth_, uint32_t depth_lod, const OrcImage* out_, uint64_t layer_stride, int pattern_step) {
  const Image& out = *(const Image*)out_;
  View input_depth(*(const Image*)depth_, (int)depth_lod, 1);
  const int cw = (int)out.width / 8 * 8, ch = (int)out.height / 4 * 4;                            // gtao.cpp:468
  const int pattern_mod = (1 << pattern_step) - 1;
  for (int py = 0; py < ch; py++)
    for (int px = 0; px < cw; px++) {
      float sampled_depth = texelFetch(input_depth, ivec2(px, py), 0).x;
      ivec2 out_pos(px >> pattern_step, py >> pattern_step);
      int layer = ((py & pattern_mod) << pattern_step) + (px & pattern_mod);
      Image l = layer_image(out, layer_stride, (uint32_t)layer);
      imageStore(l, 0, out_pos, vec4(sampled_depth, 0, 0, 0));
    }
  return 0;
}

extern "C" int orc_gtao_main_deinterleaved(const OrcGtaoParams* p, const OrcImage* depth_array_, uint64_t layer_stride, const OrcImage* normal_,
                                           const OrcImage* raw_, int pattern_n, uint32_t layer, int y0, int y1) {
  const Image& raw = *(const Image*)raw_;
  Image li = layer_image(*(const Image*)depth_array_, layer_stride, layer);
  View depth(li), normal(*(const Image*)normal_);
  const mat4 normal_mat(p->normal_mat);
  const Cam cam(p->fovy, p->aspect, p->znear, p->zfar);
  const int scale = 1 << pattern_n, scale_mod = scale - 1;
  const int cw = (int)raw.width / 8 * 8, ch = (int)raw.height / 4 * 4;                            // gtao.cpp:523
  const ivec2 tex_size(scale * cw, scale * ch);
  if (y0 < 0) y0 = 0;
  if (y1 <= 0 || y1 > ch) y1 = ch;
#pragma omp parallel for schedule(dynamic, 4)
  for (int iy = y0; iy < y1; iy++)
    for (int ix = 0; ix < cw; ix++) {
      ivec2 pixel_pos(scale * ix + (int)(layer & scale_mod), scale * iy + (int)((layer >> pattern_n) & scale_mod));
      if (pixel_pos.x >= (int)raw.width || pixel_pos.y >= (int)raw.height) continue;             // the store would be dropped: no observable effect
      vec2 screen_uv = to_vec2(pixel_pos) / to_vec2(tex_size);
      float occlusion = gtao_camera_space(depth, normal, normal_mat, cam, p->angle_offset, pixel_pos, screen_uv);
      imageStore(raw, 0, pixel_pos, vec4(occlusion, 0.f, 0.f, 0.f));
    }
  return 0;
}

// STATIC_REPROJECT: the matrix of GTAOReprojection is not read.  gtao / prev_gtao / out are R16_SFLOAT (gtao.cpp:26-36).
extern "C" int orc_gtao_reproject(const float* fovy_aspect_znear_zfar, const OrcImage* depth_, const OrcImage* prev_depth_, uint32_t depth_lod,
                                  const OrcImage* gtao_, const OrcImage* prev_gtao_, const OrcImage* out_, int y0, int y1) {
  const Image& out = *(const Image*)out_;
  View current_depth(*(const Image*)depth_, (int)depth_lod, 1), prev_depth(*(const Image*)prev_depth_, (int)depth_lod, 1);
  View current_ao(*(const Image*)gtao_), prev_ao(*(const Image*)prev_gtao_);
  const Cam cam(fovy_aspect_znear_zfar[0], fovy_aspect_znear_zfar[1], fovy_aspect_znear_zfar[2], fovy_aspect_znear_zfar[3]);
  const float REPROJECT_BIAS = 1e-6f, REPROJECT_COEF = 0.05f;
  const int cw = (int)out.width / 8 * 8, ch = (int)out.height / 4 * 4;                            // gtao.cpp:282
  const ivec2 tex_size(cw, ch);
  if (y0 < 0) y0 = 0;
  if (y1 <= 0 || y1 > ch) y1 = ch;
#pragma omp parallel for schedule(dynamic, 8)
  for (int py = y0; py < y1; py++)
    for (int px = 0; px < cw; px++) {
      ivec2 pixel_pos(px, py);
      vec2 screen_uv = to_vec2(pixel_pos) / to_vec2(tex_size);
      float new_ao = texelFetch(current_ao, pixel_pos, 0).x;
      float cur_d = texelFetch(current_depth, pixel_pos, 0).x;
      vec3 cur_view = reconstruct_view_vec(screen_uv, cur_d, cam);
      float ao = new_ao;
      float sampled_depth = texelFetch(prev_depth, pixel_pos, 0).x;
      float sampled_ao = texelFetch(prev_ao, pixel_pos, 0).x;
      float sampled_z = linearize_depth2(sampled_depth, cam.znear, cam.zfar);
      float delta = abs(sampled_z - cur_view.z);
      if (delta < REPROJECT_BIAS && sampled_depth < 1.0f) ao = mix(sampled_ao, new_ao, REPROJECT_COEF);
      imageStore(out, 0, pixel_pos, vec4(ao, 0, 0, 0));
    }
  return 0;
}
