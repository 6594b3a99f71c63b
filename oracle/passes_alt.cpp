// ORACLE — TEST INFRASTRUCTURE ONLY (see oracle.h header).  PARITY UNPINNED.
//
// Secondary image-space passes named by the north_star but outside the reference's live frame loop:
//   SSAOPass::draw     reference src/ssao.cpp:54-96, shaders/ssao/shader.frag:21-41
//   add_ssr_pass       reference src/ssr.cpp:10-73, shaders/ssr/shader.frag:28-98 + include/screen_trace.glsl:51-100
#include <cmath>
#include "oracle.h"
#include "shader_include.hpp"

using namespace orc;
static inline void row_range(int h, int& y0, int& y1) { if (y0 < 0) y0 = 0; if (y1 <= 0 || y1 > h) y1 = h; }

// ---- SSAO ---------------------------------------------------------------------------------------------------
// UBO SSAOParams (ssao.cpp:15-22): mat4 projection; fovy, aspect, znear, zfar; vec3 samples[16] (std140: 16-byte stride).
// Output is R8_UNORM (ssao.cpp:8-11).  The depth view is mip 0 with the default LINEAR sampler (:62).
extern "C" int orc_ssao(const OrcSsaoParams* p, const OrcImage* depth_, const OrcImage* out_, int y0, int y1) {
  const Image& out = *(const Image*)out_;
  View depth(*(const Image*)depth_, 0, 1);
  const mat4 projection(p->projection);
  const Cam cam(p->fovy, p->aspect, p->znear, p->zfar);
  int W = out.width, H = out.height;
  row_range(H, y0, y1);
#pragma omp parallel for schedule(dynamic, 8)
  for (int py = y0; py < y1; py++)
    for (int px = 0; px < W; px++) {
      vec2 screen_uv = (vec2((float)px, (float)py) + vec2(0.5f, 0.5f)) / vec2((float)W, (float)H);   // fullscreen triangle interpolant
      float frag_depth = texture(depth, screen_uv).x;
      vec3 camera_pos = reconstruct_view_vec(screen_uv, frag_depth, cam);
      float sum = 0.f;
      for (uint32_t i = 0; i < 16; i++) {
        vec3 s(p->samples[i][0], p->samples[i][1], p->samples[i][2]);
        vec3 pos = camera_pos + 0.05f * s;
        vec4 ndc = projection * vec4(pos, 1);
        ndc /= ndc.w;
        vec2 sample_uv = 0.5f * vec2(ndc.x, ndc.y) + vec2(0.5f, 0.5f);
        float sample_depth = texture(depth, sample_uv).x;
        float pos_depth = ndc.z;
        sum += (pos_depth < sample_depth + 0.0000001f) ? 1.0f : 0.0f;
      }
      sum /= 16;
      imageStore(out, 0, ivec2(px, py), vec4(sum, 0, 0, 0));
    }
  return 0;
}

// ---- legacy SSR -----------------------------------------------------------------------------------------------
// The depth sampler is NEAREST, clamp-to-border in U (opaque black) and clamp-to-edge in V (ssr.cpp:22-28).
static float depth_nearest(const View& v, vec2 uv) {
  int w = v.w(0), h = v.h(0);
  int x = (int)std::floor(uv.x * (float)w), y = (int)std::floor(uv.y * (float)h);
  if (x < 0 || x >= w) return 0.0f;
  y = y < 0 ? 0 : (y > h - 1 ? h - 1 : y);
  return texelFetch(v, ivec2(x, y), 0).x;
}

static float smoothstep1(float e0, float e1, float x) {
  float t = clamp((x - e0) / (e1 - e0), 0.0f, 1.0f);
  return t * t * (3.0f - 2.0f * t);
}

// UBO SSRParams (ssr.hpp:11-17).  depth: all mips (ssr.cpp:45); frame = the lit colour image; out R8G8B8A8_UNORM (:6).
extern "C" int orc_ssr_legacy(const OrcSsrParams* p, const OrcImage* depth_, const OrcImage* normal_, const OrcImage* frame_,
                              const OrcImage* material_, const OrcImage* out_, int y0, int y1) {
  const Image& out = *(const Image*)out_;
  View depth_tex(*(const Image*)depth_), normal_tex(*(const Image*)normal_), frame_tex(*(const Image*)frame_), material_tex(*(const Image*)material_);
  const mat4 camera_normal(p->normal_mat);
  const Cam cam(p->fovy, p->aspect, p->znear, p->zfar);
  int W = out.width, H = out.height;
  row_range(H, y0, y1);
#pragma omp parallel for schedule(dynamic, 4)
  for (int py = y0; py < y1; py++)
    for (int px = 0; px < W; px++) {
      vec4 out_reflection(0, 0, 0, 0);
      vec2 screen_uv = (vec2((float)px, (float)py) + vec2(0.5f, 0.5f)) / vec2((float)W, (float)H);
      do {
        vec2 tex_size = to_vec2(textureSize(frame_tex, 0));                                       // shader.frag:29
        vec2 aligned_screen_uv = floor(screen_uv * tex_size) / tex_size + 0.5f / tex_size;
        vec3 material = texture(material_tex, screen_uv).xyz();
        const float roughness = material.y;
        float pixel_depth = depth_nearest(depth_tex, aligned_screen_uv);
        vec3 pixel_normal_world = sample_gbuffer_normal(normal_tex, aligned_screen_uv);
        vec3 pixel_normal = normalize((camera_normal * vec4(pixel_normal_world, 0)).xyz());
        vec3 view_vec = reconstruct_view_vec(aligned_screen_uv, pixel_depth, cam);
        vec3 R = reflect(view_vec, pixel_normal);
        vec3 Hh = pixel_normal;
        vec3 start = project_view_vec(view_vec + 0.0005f * pixel_normal, cam);
        vec3 pp = project_view_vec(view_vec + R, cam);
        vec3 delta = normalize(pp - start);
        if (abs(delta.z) < 0.0000001f) break;                                                      // :49-52
        float t_bound = (1 - start.z) / delta.z;
        float u_bound = max((1 - start.x) / delta.x, -start.x / delta.x);
        float v_bound = max((1 - start.y) / delta.y, -start.y / delta.y);
        t_bound = min(t_bound, min(u_bound, v_bound));
        vec3 end = start + t_bound * delta;
        bool valid_hit;
        vec3 out_ray = hierarchical_raymarch(depth_tex, start, end - start, 0, 100, valid_hit);    // :66
        if (!valid_hit) break;
        vec2 screen_size = to_vec2(textureSize(frame_tex, 0));
        vec2 dist0 = abs(out_ray.xy() - start.xy());
        vec2 min_dist = vec2(2.0f) / screen_size;
        if (dist0.x < min_dist.x && dist0.y < min_dist.y) break;                                   // :73-76
        vec3 hit_normal_world = sample_gbuffer_normal(normal_tex, out_ray.xy());
        vec3 hit_normal = (camera_normal * vec4(hit_normal_world, 0)).xyz();
        if (dot(hit_normal, R) > 0) break;                                                         // :81-84
        float hit_depth = depth_nearest(depth_tex, out_ray.xy());
        if (out_ray.z > hit_depth + 0.0001f) break;                                                // :86-90
        vec2 fov = 0.05f * vec2(screen_size.y / screen_size.x, 1);
        float bx = smoothstep1(0.0f, fov.x, out_ray.x) * (1 - smoothstep1(1 - fov.x, 1.0f, out_ray.x));
        float by = smoothstep1(0.0f, fov.y, out_ray.y) * (1 - smoothstep1(1 - fov.y, 1.0f, out_ray.y));
        float coef = bx * by;
        out_reflection = coef * texture(frame_tex, out_ray.xy()) * DistributionGGX(pixel_normal, Hh, roughness) * max(dot(pixel_normal, R), 0.0f);
      } while (false);
      imageStore(out, 0, ivec2(px, py), out_reflection);
    }
  return 0;
}
