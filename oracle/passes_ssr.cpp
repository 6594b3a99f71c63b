// ORACLE — TEST INFRASTRUCTURE ONLY (see oracle.h header).  PARITY UNPINNED.
//
// Stochastic SSR: trace / filter / blur.
// reference src/shaders/advanced_ssr/{trace.comp,filter.comp,blur.comp}; bindings and
// view windows from src/advanced_ssr.cpp:147-214,:308-369,:371-438.
#include <cmath>
#include "oracle.h"
#include "shader_include.hpp"

using namespace orc;
static inline void row_range(int h, int& y0, int& y1) { if (y0 < 0) y0 = 0; if (y1 <= 0 || y1 > h) y1 = h; }

// trace.comp:143-154 (same body as gtao/main.comp:69-80)
static vec3 get_tangent(vec3 n) {
  float max_xy = max(abs(n.x), abs(n.y));
  vec3 t;
  if (max_xy < 0.00001f) t = vec3(1, 0, 0);
  else t = vec3(n.y, -n.x, 0);
  return normalize(t);
}

// trace.comp:156-158.  fract(sin(x)*43758.5453) amplifies a 1-ulp difference in sin() into
// a different Halton index, so the hash is bit-defined (SURVEY §7 "hash-driven
// stochasticity"): fp32 dot without FMA, sin evaluated in double and rounded once to
// fp32, fp32 multiply, fract.
static float rand_hash(vec2 co) {
  float d = dot(co, vec2(12.9898f, 78.233f));
  float s = (float)std::sin((double)d);
  return fract(s * 43758.5453f);
}

// trace.comp:206-268
static vec3 hierarchical_raymarch_find_hor(const View& depth_tex, vec3 origin, vec3 direction, int most_detailed_mip,
                                           uint32_t max_traversal_intersections, bool& valid_hit, vec3 w0,
                                           vec3 camera_start, float& h, const Cam& cam) {
  const vec3 inv_direction = vec3(direction.x != 0 ? 1.0f / direction.x : MAX_T_FLOAT,
                                  direction.y != 0 ? 1.0f / direction.y : MAX_T_FLOAT,
                                  direction.z != 0 ? 1.0f / direction.z : MAX_T_FLOAT);
  int current_mip = most_detailed_mip;
  vec2 screen_size = to_vec2(textureSize(depth_tex, 0));
  vec2 current_mip_resolution = get_mip_resolution(screen_size, current_mip);
  vec2 current_mip_resolution_inv = 1.0f / current_mip_resolution;
  vec2 uv_offset = 0.005f * std::exp2((float)most_detailed_mip) / screen_size;
  uv_offset.x = direction.x < 0 ? -uv_offset.x : uv_offset.x;
  uv_offset.y = direction.y < 0 ? -uv_offset.y : uv_offset.y;
  vec2 floor_offset = vec2(direction.x < 0 ? 0 : 1, direction.y < 0 ? 0 : 1);

  float current_t;
  vec3 position;
  initial_advance_ray(origin, direction, inv_direction, current_mip_resolution, current_mip_resolution_inv, floor_offset, uv_offset, position, current_t);
  h = 0;
  uint32_t i = 0;
  while (i < max_traversal_intersections && current_mip >= most_detailed_mip) {
    vec2 current_mip_position = current_mip_resolution * position.xy();
    float surface_z = texelFetch(depth_tex, to_ivec2(current_mip_position), current_mip).x;
    bool skipped_tile = advance_ray(origin, direction, inv_direction, current_mip_position, current_mip_resolution_inv, floor_offset, uv_offset, surface_z, position, current_t);
    bool mip0sample = i < 15;

    current_mip += mip0sample ? 0 : (skipped_tile ? 1 : -1);
    current_mip_resolution *= mip0sample ? 1.0f : (skipped_tile ? 0.5f : 2.0f);
    current_mip_resolution_inv *= mip0sample ? 1.0f : (skipped_tile ? 2.0f : 0.5f);
    ++i;

    vec3 v = reconstruct_view_vec(position.xy(), surface_z, cam) - camera_start;
    if (current_mip <= 1) {
      float h2 = dot(w0, normalize(v));
      if (length(v) < 0.3f)
        h = max(h, h2);
    }
  }
  valid_hit = (i <= max_traversal_intersections);
  return position;
}

extern "C" int orc_sssr_trace(const OrcTraceParams* p, const OrcImage* depth_, const OrcImage* dnormal_,
                              const OrcImage* material_, const float* halton_vec, const OrcImage* pdf_,
                              const OrcImage* out_rays_, const OrcImage* out_occl_, int y0, int y1) {
  const Image& depth_im = *(const Image*)depth_; const Image& OUT_RAY = *(const Image*)out_rays_; const Image& OUT_OCCLUSION = *(const Image*)out_occl_;
  View DEPTH(depth_im, 1, (int)depth_im.mip_count - 1);       // advanced_ssr.cpp:186 — mips 1..L-1
  View NORMAL(*(const Image*)dnormal_), MATERIAL(*(const Image*)material_), PDF_TEX(*(const Image*)pdf_);
  const mat4 normal_mat(p->normal_mat);
  const Cam cam(p->fovy, p->aspect, p->znear, p->zfar);
  const float znear = p->znear, zfar = p->zfar, max_roughness = p->max_roughness;
  const uint32_t frame_random = p->frame_random;
  const int HALTON_SEQ_SIZE = 128;
  ivec2 tex_size((int)OUT_RAY.width, (int)OUT_RAY.height);
  row_range(tex_size.y, y0, y1);

#pragma omp parallel for schedule(dynamic, 2)
  for (int py = y0; py < y1; py++)
    for (int px = 0; px < tex_size.x; px++) {
      ivec2 pixel_pos(px, py);
      vec2 screen_uv = (to_vec2(pixel_pos) + vec2(0.5f, 0.5f)) / to_vec2(tex_size);                 // :44
      vec3 material = texture(MATERIAL, screen_uv).xyz();                                             // :49

      float roughness = material.y;
      material.y = mix(0.0f, max_roughness, roughness);                                               // :52
      roughness = material.y * material.y;

      float pixel_depth = texture(DEPTH, screen_uv).x;
      vec3 pixel_normal_world = sample_gbuffer_normal(NORMAL, screen_uv);
      vec3 pixel_normal = normalize((normal_mat * vec4(pixel_normal_world, 0)).xyz());
      vec3 view_vec = reconstruct_view_vec(screen_uv, pixel_depth, cam);

      const uint32_t base_index = (uint32_t)(rand_hash(screen_uv) * HALTON_SEQ_SIZE);                 // :61
      uint32_t index = (base_index + frame_random) & (HALTON_SEQ_SIZE - 1);
      vec2 rnd(halton_vec[4 * index], halton_vec[4 * index + 1]);
      vec3 tangent = get_tangent(pixel_normal);
      vec3 bitangent = normalize(cross(pixel_normal, tangent));
      tangent = normalize(cross(bitangent, pixel_normal));

      vec3 view_dir = -normalize(view_vec);
      view_dir = vec3(dot(view_dir, tangent), dot(view_dir, bitangent), dot(view_dir, pixel_normal));

      vec3 brdf_norm = sampleGGXVNDF(view_dir, roughness, roughness, rnd.x, rnd.y);
      vec3 N = brdf_norm.x * tangent + brdf_norm.y * bitangent + brdf_norm.z * pixel_normal;
      vec3 R = reflect(view_vec, N);

      vec3 ray_start = project_view_vec(view_vec + 0.001f * pixel_normal, cam);                        // :79
      ray_start.z -= 0.0001f;

      vec3 ray_dir = project_view_vec(view_vec + R, cam);
      ray_dir -= ray_start;
      ray_dir *= (1 - ray_start.z) / ray_dir.z;

      bool valid_hit = false;
      float h = -1.0f;
      vec3 w0 = -normalize(view_vec);

      vec3 out_ray = hierarchical_raymarch_find_hor(DEPTH, ray_start, ray_dir, 0, 80, valid_hit, pixel_normal, view_vec, h, cam);  // :91
      vec4 out_ray_info = vec4(0, 0, 1, 1);

      if (valid_hit) {
        vec2 ray_step = abs(out_ray.xy() - ray_start.xy()) * to_vec2(tex_size);
        if (max(ray_step.x, ray_step.y) < 2.0f) valid_hit = false;
      }
      if (valid_hit) {
        vec3 hit_normal_world = sample_gbuffer_normal(NORMAL, out_ray.xy());
        vec3 hit_normal = (normal_mat * vec4(hit_normal_world, 0)).xyz();
        if (dot(hit_normal, R) > 0 || dot(pixel_normal, R) < 0) valid_hit = false;
      }
      if (valid_hit) {
        float hit_depth = textureLod(DEPTH, out_ray.xy(), 0).x;
        float hit_z = linearize_depth2(hit_depth, znear, zfar);
        float ray_z = linearize_depth2(out_ray.z, znear, zfar);
        if (ray_z > hit_z + 0.3f || ray_z < hit_z - 0.1f) valid_hit = false;
      }

      out_ray_info = vec4(out_ray, valid_hit ? pixel_depth : 1.0f);                                   // :120
      imageStore(OUT_RAY, 0, pixel_pos, out_ray_info);

      {
        vec3 slice_normal = normalize(cross(w0, R));
        vec3 normal_projected = pixel_normal - dot(pixel_normal, slice_normal) * slice_normal;
        vec3 X = normalize(cross(slice_normal, w0));
        float n = PI / 2.0f - std::acos(dot(normalize(normal_projected), X));

        bool no_occlusion = h == -1.0f;
        h = std::acos(h);
        h = min(n + min(h - n, PI / 2.0f), h);

        float pdf = sampleGGXdirPDF(PDF_TEX, w0, pixel_normal, R, roughness);
        float occlusion = (1.0f / PI) * length(normal_projected) * 0.25f * max(-std::cos(2 * h - n) + std::cos(n) + 2 * h * std::sin(n), 0.0f);
        float result = occlusion;
        result = std::isnan(result) ? 0.0f : result;
        imageStore(OUT_OCCLUSION, 0, pixel_pos, vec4(no_occlusion ? 0.0f : result, no_occlusion ? 0 : pdf, 0, 0));  // :138
      }
    }
  return 0;
}

// ---- filter.comp ------------------------------------------------------------
static const uint32_t NORMALIZE_REFLECTIONS = 1, BILATERAL_FILTER = 4;

static vec3 ray_weight(vec3 N, vec3 V, vec3 L, vec3 F0, float roughness) {                             // filter.comp:97-108
  vec3 H = normalize(V + L);
  vec3 F = fresnelSchlick(max(dot(H, V), 0.0f), F0);
  float alpha2 = roughness * roughness;
  float NdotL = max(dot(N, L), 0.0f);
  float NdotV = max(dot(N, V), 0.0f);
  float G2 = brdfG2(NdotL, NdotV, alpha2);
  float G1 = brdfG1(NdotV, alpha2);   // argument order as written in the reference (App. A.6 #4)
  return F * G2 / G1;
}

extern "C" int orc_sssr_filter(const OrcTraceParams* p, const OrcImage* rays_, const OrcImage* depth_, const OrcImage* albedo_,
                               const OrcImage* normal_, const OrcImage* material_, const OrcImage* out_, int y0, int y1) {
  const Image& OUT = *(const Image*)out_; const Image& depth_im = *(const Image*)depth_;
  View RAYS_TEX(*(const Image*)rays_), DEPTH_TEX(depth_im, 0, 10);                                   // advanced_ssr.cpp:342 (clamped to real mips)
  View ALBEDO_TEX(*(const Image*)albedo_), NORMAL_TEX(*(const Image*)normal_), MATERIAL_TEX(*(const Image*)material_);
  const mat4 normal_mat(p->normal_mat);
  const Cam cam(p->fovy, p->aspect, p->znear, p->zfar);
  const uint32_t render_flags = p->render_flags;
  ivec2 tex_size((int)OUT.width, (int)OUT.height);
  row_range(tex_size.y, y0, y1);
  const ivec2 offsets[5] = {ivec2(0, 0), ivec2(-1, 0), ivec2(0, 1), ivec2(1, 0), ivec2(0, -1)};

#pragma omp parallel for schedule(dynamic, 4)
  for (int py = y0; py < y1; py++)
    for (int px = 0; px < tex_size.x; px++) {
      ivec2 pixel_pos(px, py);
      vec2 screen_uv = to_vec2(pixel_pos) / to_vec2(tex_size);                                        // :39 (no +0.5)
      vec4 material = texture(MATERIAL_TEX, screen_uv);
      const float metallic = material.z;
      const float roughness = material.y;
      vec3 albedo = texture(ALBEDO_TEX, screen_uv).xyz();
      vec3 F0 = F0_approximation(albedo, metallic);
      vec3 color_sum = vec3(0);
      vec3 weight_sum = vec3(0);
      float center_depth = texelFetch(DEPTH_TEX, pixel_pos, 1).x;                                     // :75

      auto process_pixel = [&](ivec2 pos) {                                                           // :110-149
        vec4 trace_result = texelFetch(RAYS_TEX, pos, 0);
        vec2 pixel_uv = to_vec2(pos) / to_vec2(tex_size);
        float pixel_depth = texelFetch(DEPTH_TEX, pos, 1).x;
        vec3 view_vec = reconstruct_view_vec(pixel_uv, pixel_depth, cam);
        vec3 pixel_normal = sample_gbuffer_normal(NORMAL_TEX, pixel_uv);
        pixel_normal = (normal_mat * vec4(pixel_normal, 0)).xyz();
        vec3 hit_vec = reconstruct_view_vec(trace_result.xy(), trace_result.z, cam);
        vec3 radiance = (trace_result.w != 1.0f) ? texture(ALBEDO_TEX, trace_result.xy()).xyz() : vec3(0);
        vec3 V = -normalize(view_vec);
        vec3 N = pixel_normal;
        vec3 L = normalize(hit_vec - view_vec);
        vec3 weight = ray_weight(N, V, L, F0, roughness);
        float bilateral_weight = 1.0f;
        if ((render_flags & BILATERAL_FILTER) != 0)
          bilateral_weight = max(1 - 1000 * abs(center_depth - pixel_depth) / center_depth, 0.0f);
        weight *= bilateral_weight;
        color_sum += weight * radiance;
        weight_sum += weight;
      };

      if ((render_flags & NORMALIZE_REFLECTIONS) != 0) {
        for (int i = 0; i < 5; i++) process_pixel(pixel_pos + offsets[i]);
      } else {
        process_pixel(pixel_pos);
      }
      if (max(weight_sum.x, max(weight_sum.y, weight_sum.z)) < 0.001f) weight_sum = vec3(1, 1, 1);
      color_sum /= weight_sum;
      imageStore(OUT, 0, pixel_pos, vec4(color_sum, 0));
    }
  return 0;
}

// ---- blur.comp --------------------------------------------------------------
static vec3 reconstruct_world_pos_lod1(const View& depth_tex, const mat4& inverse_camera, vec2 screen_uv, const Cam& cam) {  // blur.comp:110-115
  float d = textureLod(depth_tex, screen_uv, 1).x;
  vec3 v_camera = reconstruct_view_vec(screen_uv, d, cam);
  vec4 v_world = inverse_camera * vec4(v_camera, 1.0f);
  return v_world.xyz();
}

extern "C" int orc_sssr_blur(const OrcBlurParams* p, const OrcImage* depth_, const OrcImage* normal_, const OrcImage* refl_,
                             const OrcImage* material_, const OrcImage* history_, const OrcImage* dvel_,
                             const OrcImage* prev_depth_, const OrcImage* out_, int y0, int y1) {
  const Image& OUT = *(const Image*)out_;
  View DEPTH_TEX(*(const Image*)depth_, 0, 10), HISTORY_DEPTH_TEX(*(const Image*)prev_depth_, 0, 10);  // advanced_ssr.cpp:405,:411
  View NORMAL_TEX(*(const Image*)normal_), REFLECTIONS_TEX(*(const Image*)refl_), MATERIAL_TEX(*(const Image*)material_);
  View HISTORY_TEX(*(const Image*)history_), VELOCITY_TEX(*(const Image*)dvel_);
  const mat4 inverse_camera(p->inverse_camera), prev_inverse_camera(p->prev_inverse_camera);
  const Cam cam(p->fovy_aspect_znear_zfar[0], p->fovy_aspect_znear_zfar[1], p->fovy_aspect_znear_zfar[2], p->fovy_aspect_znear_zfar[3]);
  ivec2 tex_size((int)OUT.width, (int)OUT.height);
  row_range(tex_size.y, y0, y1);

#pragma omp parallel for schedule(dynamic, 1)
  for (int py = y0; py < y1; py++)
    for (int px = 0; px < tex_size.x; px++) {
      ivec2 pixel_pos(px, py);
      vec2 screen_uv = (to_vec2(pixel_pos) + vec2(0.5f, 0.5f)) / to_vec2(tex_size);
      float roughness = texture(MATERIAL_TEX, screen_uv).y;
      roughness = mix(0.0f, p->max_roughness, roughness);
      float center_depth = texelFetch(DEPTH_TEX, pixel_pos, 1).x;
      vec3 center_normal = sample_gbuffer_normal(NORMAL_TEX, screen_uv);

      float sigma = mix(0.4f, 4, roughness);
      if (p->disable_blur != 0) sigma = 0.35f;
      float weight_sum = 0.f;
      vec3 color = vec3(0, 0, 0);
      int r = int(floor(3 * sigma - 0.01f));
      float g = 1.0f / (2 * PI * sigma * sigma);
      float e = 2 * sigma * sigma;

      for (int i = -r; i <= r; i++) {
        for (int j = -r; j <= r; j++) {
          ivec2 pos = pixel_pos + ivec2(i, j);
          vec2 uv = to_vec2(pos) / to_vec2(tex_size);
          float pixel_depth = texelFetch(DEPTH_TEX, pos, 1).x;
          vec3 pixel_normal = sample_gbuffer_normal(NORMAL_TEX, uv);
          float bilateral_weight = max(1 - 1000 * abs(center_depth - pixel_depth) / center_depth, 0.0f);
          float normal_weight = max(dot(center_normal, pixel_normal), 0.0f);
          float w = g * std::exp(-(i * i + j * j) / e);
          w *= bilateral_weight;
          w *= normal_weight;
          color += texelFetch(REFLECTIONS_TEX, pixel_pos + ivec2(i, j), 0).xyz() * w;
          weight_sum += w;
        }
      }
      color /= max(weight_sum, 0.001f);

      bool reprojected = false;
      vec2 velocity = texture(VELOCITY_TEX, screen_uv).xy();
      float delta_len = length(velocity);
      vec2 prev_uv = screen_uv + velocity;
      if (prev_uv.x >= 0 && prev_uv.y >= 0 && prev_uv.x <= 1 && prev_uv.y <= 1) {
        vec3 v_world_cur = reconstruct_world_pos_lod1(DEPTH_TEX, inverse_camera, screen_uv, cam);
        vec3 v_world_prev = reconstruct_world_pos_lod1(HISTORY_DEPTH_TEX, prev_inverse_camera, prev_uv, cam);
        vec3 v_camera = (inverse_camera * vec4(0, 0, 0, 1)).xyz();
        const float MAX_REPROJECTION_EPS = 0.1f;
        const float MIN_REPROJECTION_EPS = 0.01f;
        float error = length(v_world_cur - v_world_prev);
        float pixel_dist = length(v_world_cur - v_camera);
        float velocity_len = length(velocity);
        reprojected = (velocity_len < 0.0001f) || (error < clamp(0.1f * pixel_dist * delta_len, MIN_REPROJECTION_EPS, MAX_REPROJECTION_EPS));
      }
      if (p->accumulate == 0) reprojected = false;
      if (reprojected) {
        vec3 history_color = texture(HISTORY_TEX, screen_uv).xyz();       // :103 — screen_uv, not prev_uv (App. A.6 #3)
        color = mix(history_color, color, 0.1f);
      }
      imageStore(OUT, 0, pixel_pos, vec4(color, 0));
    }
  return 0;
}
