// ORACLE — TEST INFRASTRUCTURE ONLY (see oracle.h header).  PARITY UNPINNED.
//
// The tile path of the stochastic SSR that the reference keeps in its tree but leaves out of AdvancedSSR::run
// (src/advanced_ssr.cpp:547-550):
//   SSSR_Clear            src/advanced_ssr.cpp:440-452
//   SSSR_Classification   src/advanced_ssr.cpp:454-495, shaders/advanced_ssr/classification.comp:36-91
//   SSSR_trace (indirect) src/advanced_ssr.cpp:216-302, shaders/advanced_ssr/trace_indirect.comp:44-134
//                         (+ include/screen_trace.glsl:51-100)
//   SSSR_Tile_Regression  src/advanced_ssr.cpp:497-538, shaders/advanced_ssr/regression.comp:22-101
//
// Where the reference leaves the result open, this file fixes it and the CUDA kernels follow:
//   * the tile lists are filled through atomicAdd, so their ORDER is whatever the GPU scheduled; here tiles are
//     appended in tile-index order and every comparison is on the sorted list (the consumer writes one pixel block per
//     tile, so the order never reaches an image);
//   * shared-memory tree sums add in the order of the shader's loop (offset 32, 16, .. 1), unfused;
//   * inverse(mat3) is the cofactor form (adjugate times 1/det), the way glm and the common drivers expand it.
#include <cmath>
#include <cstring>
#include "oracle.h"
#include "shader_include.hpp"

using namespace orc;

namespace {
constexpr int TILE_SIZE = 8;
constexpr int NUM_ELEMS = TILE_SIZE * TILE_SIZE;

// classification.comp:57-76 / regression.comp:50-62: g[t] += g[t + offset] for t < offset, offset = 32 .. 1
template <typename T>
void tree_sum(T* g) {
  for (int offset = NUM_ELEMS / 2; offset != 0; offset /= 2)
    for (int t = 0; t < offset; t++) g[t] = g[t] + g[t + offset];
}

// trace_indirect.comp:136-147 (same body as trace.comp:143-154)
vec3 get_tangent(vec3 n) {
  float max_xy = max(abs(n.x), abs(n.y));
  vec3 t;
  if (max_xy < 0.00001f) t = vec3(1, 0, 0);
  else t = vec3(n.y, -n.x, 0);
  return normalize(t);
}

// trace_indirect.comp:149-151, bit-defined like oracle/passes_ssr.cpp rand_hash
float rand_hash(vec2 co) {
  float d = dot(co, vec2(12.9898f, 78.233f));
  float s = (float)std::sin((double)d);
  return fract(s * 43758.5453f);
}
}  // namespace

// classification.comp.  tex_w x tex_h = extent of `rays` (half-res); MATERIAL is the full-res image sampled at
// pixel / TEX_SIZE (no half-texel offset, App. A.6 #5).  indirect = VkDispatchIndirectCommand {x, y, z}: only .x is
// incremented.  Returns the tiles in ascending tile index.
extern "C" int orc_sssr_classification(int tex_w, int tex_h, float g_max_roughness, float g_glossy_value, const OrcImage* material_,
                                       int32_t* reflective_tiles, int32_t* glossy_tiles, uint32_t* reflective_indirect, uint32_t* glossy_indirect) {
  View MATERIAL_TEX(*(const Image*)material_);
  const int tiles_x = (tex_w + TILE_SIZE - 1) / TILE_SIZE, tiles_y = (tex_h + TILE_SIZE - 1) / TILE_SIZE;
  for (int ty = 0; ty < tiles_y; ty++)
    for (int tx = 0; tx < tiles_x; tx++) {
      float g_roughness[NUM_ELEMS];
      for (int ly = 0; ly < TILE_SIZE; ly++)
        for (int lx = 0; lx < TILE_SIZE; lx++) {
          ivec2 pixel_pos(tx * TILE_SIZE + lx, ty * TILE_SIZE + ly);
          vec2 screen_uv = to_vec2(pixel_pos) / vec2((float)tex_w, (float)tex_h);                       // :39
          float sampled_roughness = 1.f;
          if (pixel_pos.x < tex_w && pixel_pos.y < tex_h) sampled_roughness = texture(MATERIAL_TEX, screen_uv).y;
          sampled_roughness = mix(0.0f, g_max_roughness, sampled_roughness);                             // :49
          g_roughness[ly * TILE_SIZE + lx] = sampled_roughness;
        }
      tree_sum(g_roughness);
      float average_roughness = g_roughness[0] / (float)NUM_ELEMS;                                       // :79
      int tile_index = ty * tiles_x + tx;
      if (average_roughness < g_glossy_value) reflective_tiles[reflective_indirect[0]++] = tile_index;
      else glossy_tiles[glossy_indirect[0]++] = tile_index;
    }
  return 0;
}

// trace_indirect.comp.  One 8x8 block of OUT_RAY per listed tile; reflection_type 0 = mirror (march from view lod 0,
// 50 cells, plus the thickness test), 1 = glossy (view lod 1, 25 cells).  DEPTH view = mips 1..L-1 (advanced_ssr.cpp:254).
extern "C" int orc_sssr_trace_indirect(const OrcTraceParams* p, uint32_t reflection_type, const OrcImage* depth_, const OrcImage* dnormal_,
                                       const OrcImage* material_, const float* halton_vec, const int32_t* g_tiles, uint32_t tile_count,
                                       const OrcImage* out_rays_) {
  const Image& depth_im = *(const Image*)depth_;
  const Image& OUT_RAY = *(const Image*)out_rays_;
  View DEPTH(depth_im, 1, (int)depth_im.mip_count - 1);
  View NORMAL(*(const Image*)dnormal_), MATERIAL(*(const Image*)material_);
  const mat4 normal_mat(p->normal_mat);
  const Cam cam(p->fovy, p->aspect, p->znear, p->zfar);
  const float znear = p->znear, zfar = p->zfar, max_roughness = p->max_roughness;
  const uint32_t frame_random = p->frame_random;
  const int HALTON_SEQ_SIZE = 128;
  const ivec2 tex_size((int)OUT_RAY.width, (int)OUT_RAY.height);
  const int tile_width = (tex_size.x + TILE_SIZE - 1) / TILE_SIZE;                                       // :48

#pragma omp parallel for schedule(dynamic, 4)
  for (int64_t wg = 0; wg < (int64_t)tile_count; wg++) {
    int tile_index = g_tiles[wg];
    for (int ly = 0; ly < TILE_SIZE; ly++)
      for (int lx = 0; lx < TILE_SIZE; lx++) {
        ivec2 pixel_pos(tile_index % tile_width, tile_index / tile_width);
        pixel_pos = ivec2(TILE_SIZE * pixel_pos.x + lx, TILE_SIZE * pixel_pos.y + ly);
        vec2 screen_uv = to_vec2(pixel_pos) / to_vec2(tex_size);                                         // :53
        if (pixel_pos.x >= tex_size.x || pixel_pos.y >= tex_size.y) continue;

        vec3 material = texture(MATERIAL, screen_uv).xyz();
        float roughness = material.y;
        roughness = mix(0.0f, max_roughness, roughness);
        roughness *= roughness;

        float pixel_depth = texture(DEPTH, screen_uv).x;
        vec3 pixel_normal_world = sample_gbuffer_normal(NORMAL, screen_uv);
        vec3 pixel_normal = normalize((normal_mat * vec4(pixel_normal_world, 0)).xyz());
        vec3 view_vec = reconstruct_view_vec(screen_uv, pixel_depth, cam);

        const uint32_t base_index = (uint32_t)(rand_hash(screen_uv) * HALTON_SEQ_SIZE);                   // :71
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

        vec3 ray_start = project_view_vec(view_vec + 0.001f * pixel_normal, cam);                         // :89
        ray_start.z -= 0.0001f;
        vec3 ray_dir = project_view_vec(view_vec + R, cam);
        ray_dir -= ray_start;
        ray_dir *= (1 - ray_start.z) / ray_dir.z;

        bool valid_hit = false;
        vec3 out_ray(0, 0, 0);
        if (reflection_type == 0) out_ray = hierarchical_raymarch(DEPTH, ray_start, ray_dir, 0, 50, valid_hit);   // :99-103
        else out_ray = hierarchical_raymarch(DEPTH, ray_start, ray_dir, 1, 25, valid_hit);

        if (valid_hit) {                                                                                  // :107-112
          vec2 ray_step = abs(out_ray.xy() - ray_start.xy()) * to_vec2(tex_size);
          if (max(ray_step.x, ray_step.y) < 2.0f) valid_hit = false;
        }
        if (valid_hit) {                                                                                  // :114-122
          vec3 hit_normal_world = sample_gbuffer_normal(NORMAL, out_ray.xy());
          vec3 hit_normal = (normal_mat * vec4(hit_normal_world, 0)).xyz();
          if (dot(hit_normal, R) > 0 || dot(pixel_normal, R) < 0) valid_hit = false;
        }
        if (valid_hit && reflection_type == 0) {                                                          // :124-131
          float hit_depth = textureLod(DEPTH, out_ray.xy(), 0).x;
          float hit_z = linearize_depth2(hit_depth, znear, zfar);
          float ray_z = linearize_depth2(out_ray.z, znear, zfar);
          if (ray_z > hit_z + 0.3f || ray_z < hit_z - 0.1f) valid_hit = false;
        }
        imageStore(OUT_RAY, 0, pixel_pos, vec4(out_ray, valid_hit ? pixel_depth : 1.0f));                  // :133-134
      }
  }
  return 0;
}

// regression.comp.  DEPTH_TEX view = mip 1 only (advanced_ssr.cpp:524); one plane (a, b, c) with a*x + b*y + c*z = 1 fitted
// to the 64 camera-relative world positions of a half-res 8x8 tile, plus the mean squared residual in .w.  Groups cover
// ceil(w/8) x ceil(h/8); stores outside OUT_PLANES (floor(W/16) x floor(H/16)) are dropped like any imageStore.
extern "C" int orc_sssr_tile_regression(const float* camera_to_world16, const float* fovy_aspect_znear_zfar, const OrcImage* depth_,
                                        const OrcImage* out_planes_) {
  const Image& depth_im = *(const Image*)depth_;
  const Image& OUT_PLANES = *(const Image*)out_planes_;
  View DEPTH_TEX(depth_im, 1, 1);
  const mat4 camera_to_world(camera_to_world16);
  const Cam cam(fovy_aspect_znear_zfar[0], fovy_aspect_znear_zfar[1], fovy_aspect_znear_zfar[2], fovy_aspect_znear_zfar[3]);
  const ivec2 input_size = textureSize(DEPTH_TEX, 0);
  const int groups_x = (input_size.x + 7) / 8, groups_y = (input_size.y + 7) / 8;                        // advanced_ssr.cpp:536

#pragma omp parallel for schedule(dynamic, 4)
  for (int gy = 0; gy < groups_y; gy++)
    for (int gx = 0; gx < groups_x; gx++) {
      vec3 view_vecs[NUM_ELEMS], g_vec0[NUM_ELEMS], g_vec1[NUM_ELEMS], g_vec2[NUM_ELEMS];
      for (int t = 0; t < NUM_ELEMS; t++) {
        const ivec2 pixel_pos(gx * TILE_SIZE + t % TILE_SIZE, gy * TILE_SIZE + t / TILE_SIZE);
        const vec2 screen_uv = to_vec2(pixel_pos) / to_vec2(input_size);                                 // :25
        vec3 view_vec(0, 0, 0);
        if (pixel_pos.x < input_size.x && pixel_pos.y < input_size.y) {
          float pixel_depth = texelFetch(DEPTH_TEX, pixel_pos, 0).x;
          view_vec = reconstruct_view_vec(screen_uv, pixel_depth, cam);
          vec4 world_vec = camera_to_world * vec4(view_vec, 1);
          vec4 world_origin = camera_to_world * vec4(0, 0, 0, 1);
          view_vec = world_vec.xyz() - world_origin.xyz();
        }
        view_vecs[t] = view_vec;
        g_vec0[t] = view_vec;                                                                            // :41-44
        g_vec1[t] = view_vec * view_vec;
        g_vec2[t] = vec3(view_vec.x * view_vec.y, view_vec.x * view_vec.z, view_vec.y * view_vec.z);     // .xxy * .yzz
      }
      tree_sum(g_vec0); tree_sum(g_vec1); tree_sum(g_vec2);

      // :64-72  equ (symmetric) = [[s1.x s2.x s2.y] [s2.x s1.y s2.z] [s2.y s2.z s1.z]];  plane = inverse(equ) * g_vec0[0]
      const float a = g_vec1[0].x, b = g_vec2[0].x, c = g_vec2[0].y, d = g_vec1[0].y, e = g_vec2[0].z, f = g_vec1[0].z;
      const float c00 = d * f - e * e, c01 = c * e - b * f, c02 = b * e - c * d;
      const float c11 = a * f - c * c, c12 = b * c - a * e, c22 = a * d - b * b;
      const float det = a * c00 + b * c01 + c * c02;
      const float inv_det = 1.0f / det;
      const vec3 s = g_vec0[0];
      const vec3 plane((c00 * inv_det) * s.x + (c01 * inv_det) * s.y + (c02 * inv_det) * s.z,
                       (c01 * inv_det) * s.x + (c11 * inv_det) * s.y + (c12 * inv_det) * s.z,
                       (c02 * inv_det) * s.x + (c12 * inv_det) * s.y + (c22 * inv_det) * s.z);

      float sse_v[NUM_ELEMS];
      for (int t = 0; t < NUM_ELEMS; t++) {                                                              // :76-85
        float sse = dot(plane, view_vecs[t]) - 1;
        sse *= sse;
        if (std::isnan(sse)) sse = 1e10f;
        sse_v[t] = sse;
      }
      tree_sum(sse_v);
      imageStore(OUT_PLANES, 0, ivec2(gx, gy), vec4(plane, sse_v[0] / (float)NUM_ELEMS));                 // :98-101
    }
  return 0;
}
