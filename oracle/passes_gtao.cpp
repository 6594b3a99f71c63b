// ORACLE — TEST INFRASTRUCTURE ONLY (see oracle.h header).  PARITY UNPINNED.
//
// GTAO main (MIS) / filter / temporal accumulate.
// reference src/shaders/gtao/{main.comp,filter.comp,accum.comp}; bindings from
// src/gtao.cpp:84-148,:198-239,:286-347.
#include <cmath>
#include "oracle.h"
#include "shader_include.hpp"

using namespace orc;
static inline void row_range(int h, int& y0, int& y1) { if (y0 < 0) y0 = 0; if (y1 <= 0 || y1 > h) y1 = h; }

static const float MAX_THIKNESS = 0.1f;                                                              // main.comp:82

// main.comp:84-108
static float find_horizon(const View& depth, vec2 start, vec3 camera_start, vec2 dir, int samples_count, vec3 v, const Cam& cam) {
  float h_cos = -1.0f;
  float previous_z = camera_start.z;
  for (int i = 1; i <= samples_count; i++) {
    vec2 tc = start + (float(i) / samples_count) * dir;
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

// main.comp:276-278
static float gtao_direction(ivec2 pos) { return (1.0f / 16.0f) * ((((pos.x + pos.y) & 3) << 2) + (pos.x & 3)); }

struct GtaoCtx {
  View depth, gbuffer_normal, gbuffer_material, pdf;
  const Image* gtao_out;
  mat4 normal_mat; Cam cam;
  float angle_offset, weight_ratio; uint32_t reflections_only;
};

// main.comp:185-217
static float gtao_camera_space(const GtaoCtx& c, ivec2 pos, vec2 screen_uv, uint32_t dirs_count) {
  float frag_depth = texture(c.depth, screen_uv).x;
  if (frag_depth >= 1.f) return 0.f;
  vec3 camera_pos = reconstruct_view_vec(screen_uv, frag_depth, c.cam);
  vec3 w0 = -normalize(camera_pos);
  vec3 camera_normal = normalize((c.normal_mat * vec4(decode_normal(texture(c.gbuffer_normal, screen_uv).xy()), 0)).xyz());
  vec2 dir_radius = min(100.0f / length(camera_pos), 16.0f) / to_vec2(textureSize(c.depth, 0));
  float base_angle = gtao_direction(pos) + c.angle_offset;
  float sum = 0.f;
  for (uint32_t dir_index = 0; dir_index < dirs_count; dir_index++) {
    float angle = 2 * PI * (base_angle + float(dir_index) / float(dirs_count));
    vec2 sample_direction = dir_radius * vec2(std::cos(angle), std::sin(angle));
    vec3 sample_end_pos = reconstruct_view_vec(screen_uv + sample_direction, frag_depth, c.cam);
    vec3 slice_normal = normalize(cross(w0, -sample_end_pos));
    vec3 normal_projected = camera_normal - dot(camera_normal, slice_normal) * slice_normal;
    vec3 X = -normalize(cross(slice_normal, w0));
    float n = PI / 2.0f - std::acos(dot(normalize(normal_projected), X));
    float h_cos = find_horizon(c.depth, screen_uv, camera_pos, sample_direction, 16, w0, c.cam);
    float h = std::acos(h_cos);
    h = min(n + min(h - n, PI / 2.0f), h);
    sum += length(normal_projected) * 0.25f * max(-std::cos(2 * h - n) + std::cos(n) + 2 * h * std::sin(n), 0.0f);
  }
  return 2 * sum / float(dirs_count);
}

// main.comp:219-274
static vec2 mis_gtao(const GtaoCtx& c, ivec2 pos, vec2 screen_uv) {
  float frag_depth = texture(c.depth, screen_uv).x;
  if (frag_depth >= 1.f) return vec2(0.f, 1.0f);

  vec3 camera_pos = reconstruct_view_vec(screen_uv, frag_depth, c.cam);
  vec3 w0 = -normalize(camera_pos);
  vec3 camera_normal = normalize((c.normal_mat * vec4(decode_normal(texture(c.gbuffer_normal, screen_uv).xy()), 0)).xyz());

  vec2 dir_radius = min(100.0f / length(camera_pos), 16.0f) / to_vec2(textureSize(c.depth, 0));
  float base_angle = gtao_direction(pos) + c.angle_offset;
  float angle = 2 * PI * base_angle;

  vec2 sample_direction = dir_radius * vec2(std::cos(angle), std::sin(angle));
  vec3 sample_end_pos = reconstruct_view_vec(screen_uv + sample_direction, frag_depth, c.cam);
  vec3 L = normalize(sample_end_pos - camera_pos);

  vec3 slice_normal = normalize(cross(w0, -sample_end_pos));
  vec3 normal_projected = camera_normal - dot(camera_normal, slice_normal) * slice_normal;
  vec3 X = -normalize(cross(slice_normal, w0));
  float n = PI / 2.0f - std::acos(dot(normalize(normal_projected), X));

  float h_cos = find_horizon(c.depth, screen_uv, camera_pos, sample_direction, 16, w0, c.cam);
  float h = std::acos(h_cos);
  h = min(n + min(h - n, PI / 2.0f), h);
  float occlusion = (1.0f / PI) * length(normal_projected) * 0.25f * max(-std::cos(2 * h - n) + std::cos(n) + 2 * h * std::sin(n), 0.0f);

  float roughness = texture(c.gbuffer_material, screen_uv).y;
  float pdf_ggx = sampleGGXdirPDF(c.pdf, w0, camera_normal, L, roughness * roughness);
  float pdf_uniform = 1.0f / (2 * PI);

  vec2 ao = imageLoad(*c.gtao_out, 0, pos).xy();

  if (c.reflections_only != 0) {
    float res = ao.x / ao.y;
    return vec2(std::isnan(res) ? 1 : res, 1);
  }

  float alpha = 1 / (c.weight_ratio + 1);
  float betta = 1 - alpha;
  float mis_weight1 = alpha / (alpha * ao.y + betta * pdf_uniform);
  float mis_weight2 = betta / (alpha * pdf_ggx + betta * pdf_uniform);
  float mis_ao = ao.x * mis_weight1 + occlusion * mis_weight2;
  float result = mis_ao;
  float total_weight = 1;
  return vec2(std::isnan(result) ? occlusion / pdf_uniform : mis_ao, total_weight);
}

extern "C" int orc_gtao_main(const OrcGtaoParams* p, const OrcImage* depth_, const OrcImage* normal_, const OrcImage* material_,
                             const OrcImage* pdf_, const OrcImage* raw_, uint32_t depth_lod, int y0, int y1) {
  const Image& raw = *(const Image*)raw_;
  GtaoCtx c{View(*(const Image*)depth_, (int)depth_lod, 1), View(*(const Image*)normal_), View(*(const Image*)material_),
            View(*(const Image*)pdf_), &raw, mat4(p->normal_mat), Cam(p->fovy, p->aspect, p->znear, p->zfar),
            p->angle_offset, p->weight_ratio, p->reflections_only};
  // gtao.cpp:145 dispatch(w/8, h/4) truncates; main.comp:54 tex_size = groups * local size.
  ivec2 tex_size((int)(raw.width / 8) * 8, (int)(raw.height / 4) * 4);
  row_range(tex_size.y, y0, y1);
#pragma omp parallel for schedule(dynamic, 4)
  for (int py = y0; py < y1; py++)
    for (int px = 0; px < tex_size.x; px++) {
      ivec2 pixel_pos(px, py);
      vec2 screen_uv = (to_vec2(pixel_pos) + vec2(0.5f, 0.5f)) / to_vec2(tex_size);
      vec2 occlusion = vec2(0, 1 / (2 * PI));
      if (p->use_mis > 0) occlusion = mis_gtao(c, pixel_pos, screen_uv);
      else occlusion.x = gtao_camera_space(c, pixel_pos, screen_uv, (p->two_directions != 0) ? 2 : 1);
      imageStore(raw, 0, pixel_pos, vec4(occlusion.x, occlusion.y, 0.f, 0.f));
    }
  return 0;
}

// filter.comp:17-52
extern "C" int orc_gtao_filter(const OrcGtaoParams* p, const OrcImage* depth_, const OrcImage* raw_, const OrcImage* filtered_,
                               uint32_t depth_lod, int y0, int y1) {
  const Image& filtered = *(const Image*)filtered_;
  View depth_tex(*(const Image*)depth_, (int)depth_lod, 1), raw_gtao(*(const Image*)raw_);
  const float znear = p->znear, zfar = p->zfar;
  ivec2 tex_size((int)(filtered.width / 8) * 8, (int)(filtered.height / 4) * 4);      // gtao.cpp:237 truncating dispatch
  row_range(tex_size.y, y0, y1);
#pragma omp parallel for schedule(dynamic, 4)
  for (int py = y0; py < y1; py++)
    for (int px = 0; px < tex_size.x; px++) {
      ivec2 pixel_pos(px, py);
      float pixel_depth = texelFetch(depth_tex, pixel_pos, 0).x;
      if (pixel_depth >= 1.f) imageStore(filtered, 0, pixel_pos, vec4(1, 0, 0, 0));    // :23-25 — falls through (App. A.6 #2)
      float linear_depth = linearize_depth2(pixel_depth, znear, zfar);
      float weight_sum = 0;
      float norm_weigth = 0;
      float ao = 0;
      for (int x = 0; x < 4; x++) {
        for (int y = 0; y < 4; y++) {
          ivec2 sample_pos = pixel_pos + ivec2(x - 2, y - 2);
          float sampled_depth = linearize_depth2(texelFetch(depth_tex, sample_pos, 0).x, znear, zfar);
          float weight = max(0.0f, 1.0f - 5 * abs(sampled_depth - linear_depth) / abs(linear_depth));
          weight_sum += weight;
          ao += weight * texelFetch(raw_gtao, sample_pos, 0).x;
          norm_weigth += weight * texelFetch(raw_gtao, sample_pos, 0).y;
        }
      }
      (void)norm_weigth;
      ao /= weight_sum;
      imageStore(filtered, 0, pixel_pos, vec4(ao, 0, 0, 0));
    }
  return 0;
}

// accum.comp:94-99
static vec3 reconstruct_world_pos(const View& depth_tex, const mat4& inverse_camera, vec2 screen_uv, const Cam& cam) {
  float d = texture(depth_tex, screen_uv).x;
  vec3 v_camera = reconstruct_view_vec(screen_uv, d, cam);
  vec4 v_world = inverse_camera * vec4(v_camera, 1.0f);
  return v_world.xyz();
}

// accum.comp:29-92
extern "C" int orc_gtao_accumulate(const OrcAccumParams* p, const OrcImage* depth_, const OrcImage* prev_depth_,
                                   const OrcImage* filtered_, const OrcImage* accumulated_, const OrcImage* dvel_,
                                   const OrcImage* history_, uint32_t depth_lod, int y0, int y1) {
  const Image& accumulated_ao = *(const Image*)accumulated_;
  View current_depth(*(const Image*)depth_, (int)depth_lod, 1), prev_depth(*(const Image*)prev_depth_, (int)depth_lod, 1);
  View current_ao(*(const Image*)filtered_), velocity_tex(*(const Image*)dvel_), accumulated_history(*(const Image*)history_);
  const mat4 inverse_camera(p->inverse_camera), prev_inverse_camera(p->prev_inverse_camera), mvp(p->mvp);
  const Cam cam(p->fovy_aspect_znear_zfar[0], p->fovy_aspect_znear_zfar[1], p->fovy_aspect_znear_zfar[2], p->fovy_aspect_znear_zfar[3]);
  const float MAX_SAMPLES = 255.f;
  ivec2 tex_size((int)accumulated_ao.width, (int)accumulated_ao.height);
  row_range(tex_size.y, y0, y1);
#pragma omp parallel for schedule(dynamic, 8)
  for (int py = y0; py < y1; py++)
    for (int px = 0; px < tex_size.x; px++) {
      ivec2 pixel_pos(px, py);
      vec2 screen_uv = (to_vec2(pixel_pos) + vec2(0.5f, 0.5f)) / to_vec2(tex_size);
      vec2 velocity = texture(velocity_tex, screen_uv).xy();
      vec2 prev_uv = screen_uv + velocity;
      bool reprojected = false;
      float valid_samples = 1.0f;
      if (prev_uv.x >= 0 && prev_uv.y >= 0 && prev_uv.x <= 1 && prev_uv.y <= 1) {
        vec3 v_world_cur = reconstruct_world_pos(current_depth, inverse_camera, screen_uv, cam);
        (void)v_world_cur;
        vec3 v_world_prev = reconstruct_world_pos(prev_depth, prev_inverse_camera, prev_uv, cam);
        vec4 prev_ndc = mvp * vec4(v_world_prev, 1.0f);
        prev_ndc /= prev_ndc.w;
        vec2 prev_world_uv = 0.5f * prev_ndc.xy() + vec2(0.5f, 0.5f);
        vec2 delta = abs(prev_world_uv - screen_uv) * to_vec2(tex_size);
        const float znear = cam.znear, zfar = cam.zfar;
        float current_z = linearize_depth2(texture(current_depth, screen_uv).x, znear, zfar);
        float prev_z = linearize_depth2(prev_ndc.z, znear, zfar);
        float depth_err = abs((prev_z - current_z));
        float vel_delta = max(abs(velocity.x) * tex_size.x, abs(velocity.y) * tex_size.y);
        float error = 0.1f * vel_delta + depth_err;
        valid_samples = clamp(1 - error, 0.8f, 1);
        reprojected = (max(delta.x, delta.y) <= 2.0f) && (depth_err < 0.2f);
      }
      float new_ao = texelFetch(current_ao, pixel_pos, 0).x;
      float computed_ao = new_ao;
      float samples_count = 1.f;
      if (p->clear_history != 0) reprojected = false;
      if (reprojected) {
        vec2 accumulated = texture(accumulated_history, prev_uv).xy();
        samples_count = 255.f * accumulated.y * valid_samples;
        computed_ao = (accumulated.x * samples_count + new_ao) / (samples_count + 1.f);
        samples_count += 1.f;
        if (samples_count > MAX_SAMPLES) samples_count = 100.f;
      }
      imageStore(accumulated_ao, 0, pixel_pos, vec4(clamp(computed_ao, 0, 1), samples_count / 255.f, 0.f, 0.f));
    }
  return 0;
}
