// ORACLE — TEST INFRASTRUCTURE ONLY (see oracle.h header).  PARITY UNPINNED.
//
// Deferred PBR lighting and TAA resolve.
// reference src/shaders/defered_shading/shader.frag:41-130, src/shaders/taa/resolve.comp:20-77;
// bindings from src/defered_shading.cpp:75-117 and src/taa.cpp:32-62.
#include <cmath>
#include "oracle.h"
#include "shader_include.hpp"

using namespace orc;
static inline void row_range(int h, int& y0, int& y1) { if (y0 < 0) y0 = 0; if (y1 <= 0 || y1 > h) y1 = h; }

// shader.frag:103-130
static vec4 sample_ocllusion_ssr(const View& depth_tex, const View& occlusion_tex, const View& reflections_tex, float depth, vec2 screen_uv) {
  vec4 lowres_depth;
  lowres_depth.x = textureLodOffset(depth_tex, screen_uv, 1, ivec2(0, 0)).x;
  lowres_depth.y = textureLodOffset(depth_tex, screen_uv, 1, ivec2(1, 0)).x;
  lowres_depth.z = textureLodOffset(depth_tex, screen_uv, 1, ivec2(0, 1)).x;
  lowres_depth.w = textureLodOffset(depth_tex, screen_uv, 1, ivec2(1, 1)).x;

  vec4 delta = abs(lowres_depth - vec4(depth, depth, depth, depth));
  float min_delta = min(min(delta.x, delta.y), min(delta.z, delta.w));
  vec4 result = vec4(0, 0, 0, 0);
  ivec2 off;
  if (min_delta == delta.x) off = ivec2(0, 0);
  else if (min_delta == delta.y) off = ivec2(1, 0);
  else if (min_delta == delta.z) off = ivec2(0, 1);
  else off = ivec2(1, 1);
  result.w = textureOffset(occlusion_tex, screen_uv, off).x;
  vec3 r = textureOffset(reflections_tex, screen_uv, off).xyz();
  result.x = r.x; result.y = r.y; result.z = r.z;
  return result;
}

extern "C" int orc_deferred_shading(const OrcShadingParams* p, const OrcImage* albedo_, const OrcImage* normal_,
                                    const OrcImage* material_, const OrcImage* depth_, const OrcImage* occlusion_,
                                    const OrcImage* brdf_, const OrcImage* reflections_, const OrcImage* out_, int y0, int y1) {
  const Image& out = *(const Image*)out_;
  View albedo_tex(*(const Image*)albedo_), normal_tex(*(const Image*)normal_), material_tex(*(const Image*)material_);
  View depth_tex(*(const Image*)depth_), occlusion_tex(*(const Image*)occlusion_), brdf_tex(*(const Image*)brdf_), reflections_tex(*(const Image*)reflections_);
  const mat4 inverse_camera(p->inverse_camera);
  const Cam cam(p->fovy, p->aspect, p->znear, p->zfar);
  const vec3 LIGHT_POS = vec3(-1.85867f, 5.81832f, -0.247114f);                                       // :37
  const vec3 LIGHT_RADIANCE = vec3(0.1f, 0.1f, 0.1f);                                                 // :38
  int W = out.width, H = out.height;
  row_range(H, y0, y1);
#pragma omp parallel for schedule(dynamic, 8)
  for (int py = y0; py < y1; py++)
    for (int px = 0; px < W; px++) {
      // fullscreen triangle interpolant (include/screen_verts.glsl:1-14) -> pixel-centre uv
      vec2 screen_uv = (vec2((float)px, (float)py) + vec2(0.5f, 0.5f)) / vec2((float)W, (float)H);
      vec3 normal = sample_gbuffer_normal(normal_tex, screen_uv);
      vec3 albedo = texture(albedo_tex, screen_uv).xyz();
      vec4 material = texture(material_tex, screen_uv);
      float depth = textureLod(depth_tex, screen_uv, 0).x;
      vec4 ssr_occlusion = sample_ocllusion_ssr(depth_tex, occlusion_tex, reflections_tex, depth, screen_uv);
      float occlusion = ssr_occlusion.w;
      vec3 reflection = ssr_occlusion.xyz();

      vec3 camera_view_vec = reconstruct_view_vec(screen_uv, depth, cam);
      vec3 world_pos = (inverse_camera * vec4(camera_view_vec, 1)).xyz();
      vec3 camera_pos = (inverse_camera * vec4(0, 0, 0, 1)).xyz();

      const float metallic = mix(0.1f, 1.0f, material.z);
      const float roughness = material.y;

      const vec3 V = normalize(camera_pos - world_pos);
      const vec3 N = normal;

      vec3 F0 = F0_approximation(albedo, metallic);
      vec3 Lo = vec3(0);

      vec3 L = normalize(LIGHT_POS - world_pos);
      vec3 Hv = normalize(V + L);

      float light_distance = length(LIGHT_POS - world_pos);
      vec3 radiance = LIGHT_RADIANCE * min(100 / (light_distance * light_distance), 100.0f);

      float NdotL = max(dot(N, L), 0.0f);
      float NdotV = max(dot(N, V), 0.0f);

      float NDF = DistributionGGX(N, Hv, roughness);
      float G = brdfG2(NdotV, NdotL, roughness * roughness);
      vec3 F = fresnelSchlick(max(dot(Hv, V), 0.0f), F0);

      vec3 kS = F;
      vec3 kD = (vec3(1.0f) - kS) * (1 - metallic);

      vec3 specular = (NDF * G * F) / (4.0f * NdotV * NdotL + 0.0001f);
      float biased_rougness = mix(p->min_max_roughness[0], p->min_max_roughness[1], roughness);
      vec2 ssr_brdf = texture(brdf_tex, vec2(biased_rougness, NdotV)).xy();

      Lo += (kD * albedo / PI + specular) * radiance * NdotL;
      Lo += reflection * (F0 * ssr_brdf.x + vec3(ssr_brdf.y));
      vec3 color = occlusion * (vec3(0.6f) * albedo + Lo);

      vec4 out_color;
      if (p->show_ao != 0) out_color = vec4(occlusion, occlusion, occlusion, 0);
      else out_color = vec4(color, 0);
      imageStore(out, 0, ivec2(px, py), out_color);
    }
  return 0;
}

// resolve.comp:72-77
static vec3 reconstruct_world_pos(const View& depth_tex, const mat4& inverse_camera, vec2 screen_uv, const Cam& cam) {
  float d = texture(depth_tex, screen_uv).x;
  vec3 v_camera = reconstruct_view_vec(screen_uv, d, cam);
  vec4 v_world = inverse_camera * vec4(v_camera, 1.0f);
  return v_world.xyz();
}

// resolve.comp:20-70
extern "C" int orc_taa_resolve(const OrcTaaParams* p, const OrcImage* history_, const OrcImage* prev_depth_, const OrcImage* depth_,
                               const OrcImage* velocity_, const OrcImage* color_, const OrcImage* out_, int y0, int y1) {
  const Image& OUTPUT = *(const Image*)out_;
  View HISTORY_COLOR_TEX(*(const Image*)history_), HISTORY_DEPTH_TEX(*(const Image*)prev_depth_), CURRENT_DEPTH_TEX(*(const Image*)depth_);
  View VELOCITY_TEX(*(const Image*)velocity_), COLOR_TEX(*(const Image*)color_);
  const mat4 inverse_camera(p->inverse_camera), prev_inverse_camera(p->prev_inverse_camera);
  const Cam cam(p->fovy_aspect_znear_zfar[0], p->fovy_aspect_znear_zfar[1], p->fovy_aspect_znear_zfar[2], p->fovy_aspect_znear_zfar[3]);
  ivec2 tex_size((int)OUTPUT.width, (int)OUTPUT.height);
  row_range(tex_size.y, y0, y1);
#pragma omp parallel for schedule(dynamic, 8)
  for (int py = y0; py < y1; py++)
    for (int px = 0; px < tex_size.x; px++) {
      ivec2 pixel_pos(px, py);
      vec2 screen_uv = (to_vec2(pixel_pos) + vec2(0.5f, 0.5f)) / to_vec2(tex_size);
      vec3 out_color = vec3(0, 0, 0);
      vec3 current_color = texture(COLOR_TEX, screen_uv).xyz();
      vec2 velocity = texture(VELOCITY_TEX, screen_uv).xy();
      float delta_len = length(velocity);
      bool reprojected = false;
      vec2 prev_uv = screen_uv + velocity;

      if (prev_uv.x >= 0 && prev_uv.y >= 0 && prev_uv.x <= 1 && prev_uv.y <= 1) {
        vec3 history = texture(HISTORY_COLOR_TEX, prev_uv).xyz();
        vec3 color0 = textureOffset(HISTORY_COLOR_TEX, prev_uv, ivec2(1, 0)).xyz();
        vec3 color1 = textureOffset(HISTORY_COLOR_TEX, prev_uv, ivec2(0, 1)).xyz();
        vec3 color2 = textureOffset(HISTORY_COLOR_TEX, prev_uv, ivec2(-1, 0)).xyz();
        vec3 color3 = textureOffset(HISTORY_COLOR_TEX, prev_uv, ivec2(0, -1)).xyz();

        vec3 color_min = min(color0, min(color1, min(color2, color3)));
        vec3 color_max = max(color0, max(color1, max(color2, color3)));
        history = clamp(history, color_min, color_max);

        out_color = mix(history, current_color, 0.1f);

        vec3 v_world_cur = reconstruct_world_pos(CURRENT_DEPTH_TEX, inverse_camera, screen_uv, cam);
        vec3 v_world_prev = reconstruct_world_pos(HISTORY_DEPTH_TEX, prev_inverse_camera, prev_uv, cam);
        vec3 v_camera = (inverse_camera * vec4(0, 0, 0, 1)).xyz();

        const float MAX_REPROJECTION_EPS = 0.2f;
        const float MIN_REPROJECTION_EPS = 0.01f;
        float error = length(v_world_cur - v_world_prev);
        float pixel_dist = length(v_world_cur - v_camera);
        reprojected = (delta_len < 0.005f) || (error < clamp(0.1f * pixel_dist * delta_len, MIN_REPROJECTION_EPS, MAX_REPROJECTION_EPS));
      }
      if (!reprojected) out_color = current_color;
      imageStore(OUTPUT, 0, pixel_pos, vec4(out_color, 0.f));
    }
  return 0;
}
