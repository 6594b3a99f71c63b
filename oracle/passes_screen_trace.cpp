// ORACLE — TEST INFRASTRUCTURE ONLY (see oracle.h header).  PARITY UNPINNED.
//
// ScreenSpaceTrace: combined AO + one-bounce colour gather in tangent space, its 4x4 depth-aware filter and the static
// temporal accumulation.
//   add_main_pass        reference src/screen_trace.cpp:23-99,   shaders/screen_trace/trace.comp:27-37,:238-343
//   add_filter_pass      reference src/screen_trace.cpp:101-139, shaders/screen_trace/filter.comp:13-39
//   add_accumulate_pass  reference src/screen_trace.cpp:141-180, shaders/screen_trace/accumulate.comp:20-40
//
// Two things the reference leaves undefined are DEFINED here (and identically in the CUDA kernel):
//   * trace.comp exchanges ray hits between the 64 invocations of a tile through shared memory with memoryBarrierShared()
//     but no barrier() (:305-306), and sky pixels return before writing their slot (:240-242).  Here every invocation of
//     the tile writes its slot first (sky pixels write "no hit"), then all of them gather: the result a barrier() would give.
//   * The dispatches cover floor(W/8) x floor(H/8) groups (trace) and floor(W/8) x floor(H/4) (filter, accumulate); pixels
//     beyond are never written, and `tex_size` in the shaders is the covered extent, not the image extent.
// The three direction angles are evaluated in double and rounded once to fp32 (like rand(), passes_ssr.cpp:22-30), so both
// implementations start every march from the same bits.
#include <cmath>
#include <vector>
#include "oracle.h"
#include "shader_include.hpp"

using namespace orc;

namespace {

const float MAX_THIKNESS = 0.2f;
const int SAMPLES = 20;
const int TILE_SIZE = 8;

float rand_glsl(vec2 co) {                                                                     // trace.comp:42-44
  float d = dot(co, vec2(12.9898f, 78.233f));
  float s = (float)std::sin((double)d);
  return fract(s * 43758.5453f);
}
float sin_r(float x) { return (float)std::sin((double)x); }
float cos_r(float x) { return (float)std::cos((double)x); }

float gtao_direction(ivec2 pos) {                                                              // :135-137
  return (1.0f / 16.0f) * (float)((((pos.x + pos.y) & 3) << 2) + (pos.x & 3));
}

void calc_tangent_space(vec3 normal, vec3& tangent, vec3& bitangent) {                         // :220-232
  if (abs(normal.z) > 0.0f) {
    float k = sqrt(normal.y * normal.y + normal.z * normal.z);
    tangent = vec3(0.0f, -normal.z / k, normal.y / k);
  } else {
    float k = sqrt(normal.x * normal.x + normal.y * normal.y);
    tangent = vec3(normal.y / k, -normal.x / k, 0.0f);
  }
  bitangent = cross(normal, tangent);
}

struct TraceCtx {
  View depth, normal_tex, color, material;
  mat4 normal_mat;
  Cam cam;
  float random_offset, angle_offset;
};

vec3 sample_normal(const TraceCtx& c, vec2 uv) {                                               // :46-48
  return normalize((c.normal_mat * vec4(decode_normal(texture(c.normal_tex, uv).xy()), 0)).xyz());
}

struct Lane {            // what one invocation carries from the march phase to the gather phase
  bool sky = true;
  vec3 screen_pos, camera_pos, camera_normal;
  float ao = 0.f;
  vec3 hit_pos = vec3(-1, -1, -1), hit_color;
};

// trace_tangent_space up to the shared-memory stores (:238-305), dirs_count = 1 (main(): :35)
void march(const TraceCtx& c, ivec2 pos, vec2 screen_uv, Lane& L) {
  L.screen_pos = vec3(screen_uv, textureLod(c.depth, screen_uv, 0).x);
  if (L.screen_pos.z >= 1.f) return;
  L.sky = false;
  vec3 screen_pos = L.screen_pos;
  vec3 camera_pos = reconstruct_view_vec(screen_pos.xy(), screen_pos.z, c.cam);
  vec3 camera_normal = sample_normal(c, screen_pos.xy());
  camera_pos += 1e-6f * camera_normal;
  vec3 tangent, bitangent;
  calc_tangent_space(camera_normal, tangent, bitangent);

  float base_angle = gtao_direction(pos) + c.angle_offset;
  float normal_angle = PI / 2 * rand_glsl(screen_uv + vec2(c.random_offset, 0));
  vec2 ao_dir_radius = min(200.0f / length(camera_pos), 32.0f) / to_vec2(textureSize(c.depth, 0));

  float angle = 2 * PI * (base_angle + 0.0f / 1.0f);
  vec3 camera_sample_dir = normalize(cos_r(angle) * tangent + sin_r(angle) * bitangent + camera_normal * sin_r(normal_angle));
  vec3 screen_dir = project_view_vec(camera_pos + camera_sample_dir, c.cam);
  screen_dir -= screen_pos;
  screen_dir = screen_dir / length(screen_dir.xy()) * max(ao_dir_radius.x, ao_dir_radius.y);
  bool ray_hit = false;
  vec3 hit_pos(0, 0, 0);

  float h_cos = 0;
  float previous_z = camera_pos.z;
  for (int i = 0; i < SAMPLES; i++) {
    vec3 tc = screen_pos + ((float)i / (float)SAMPLES) * screen_dir;
    float tc_depth = textureLod(c.depth, tc.xy(), 0).x;
    vec3 camera_sample = reconstruct_view_vec(tc.xy(), tc_depth, c.cam);
    if (tc.x < 0 || tc.x > 1 || tc.y < 0 || tc.y > 1 || camera_sample.z > previous_z + MAX_THIKNESS) break;
    if (!ray_hit && tc.z - 1e-6f > tc_depth) { hit_pos = tc; ray_hit = true; }
    h_cos = max(h_cos, dot(camera_normal, normalize(camera_sample - camera_pos)));
    previous_z = camera_sample.z;
  }
  h_cos = min(h_cos, 1.0f);
  float h = std::acos(h_cos);
  L.ao = 0.25f * (1 - std::cos(2 * h));

  const int FAR_SAMPLES = 8;                                                                    // :289-304
  vec3 start_ray = screen_pos + screen_dir;
  screen_dir *= 2.0f;
  for (int i = 0; i < FAR_SAMPLES; i++) {
    vec3 tc = start_ray + ((float)i / (float)FAR_SAMPLES) * screen_dir;
    float tc_depth = textureLod(c.depth, tc.xy(), 0).x;
    float camera_z = linearize_depth2(tc_depth, c.cam.znear, c.cam.zfar);
    if (tc.x < 0 || tc.x > 1 || tc.y < 0 || tc.y > 1 || camera_z > previous_z + 0.1f) break;
    if (!ray_hit && tc.z - 1e-6f > tc_depth) { hit_pos = tc; ray_hit = true; }
    previous_z = camera_z;
  }
  vec3 hit_normal = ray_hit ? sample_normal(c, hit_pos.xy()) : vec3(0, 0, 0);
  ray_hit = ray_hit && (dot(camera_normal, hit_normal) < 0);
  L.hit_pos = ray_hit ? hit_pos : vec3(-1, -1, -1);
  L.hit_color = ray_hit ? texture(c.color, hit_pos.xy()).xyz() : vec3(0, 0, 0);
  L.camera_pos = camera_pos;
  L.camera_normal = camera_normal;
}

}  // namespace

// Rows [y0, y1) are processed in whole 8-row tiles (tiles whose first row lies in the range).
extern "C" int orc_screen_trace_main(const OrcScreenTraceParams* p, const OrcImage* depth_, const OrcImage* normal_, const OrcImage* color_,
                                     const OrcImage* material_, const OrcImage* out_, int y0, int y1) {
  const Image& out = *(const Image*)out_;
  TraceCtx c{View(*(const Image*)depth_, 0, 1), View(*(const Image*)normal_), View(*(const Image*)color_), View(*(const Image*)material_),
             mat4(p->normal_mat), Cam(p->fovy, p->aspect, p->znear, p->zfar), p->random_offset, p->angle_offset};
  const int gx = (int)out.width / TILE_SIZE, gy = (int)out.height / TILE_SIZE;                 // cmd.dispatch(extent.width/8, extent.height/8, 1)
  const ivec2 tex_size(gx * TILE_SIZE, gy * TILE_SIZE);                                        // trace.comp:29
  if (y0 < 0) y0 = 0;
  if (y1 <= 0 || y1 > (int)out.height) y1 = (int)out.height;
#pragma omp parallel for schedule(dynamic, 1)
  for (int ty = 0; ty < gy; ty++) {
    if (ty * TILE_SIZE < y0 || ty * TILE_SIZE >= y1) continue;
    std::vector<Lane> lanes(TILE_SIZE * TILE_SIZE);
    for (int tx = 0; tx < gx; tx++) {
      for (int ly = 0; ly < TILE_SIZE; ly++)
        for (int lx = 0; lx < TILE_SIZE; lx++) {
          ivec2 pixel_pos(tx * TILE_SIZE + lx, ty * TILE_SIZE + ly);
          vec2 screen_uv = to_vec2(pixel_pos) / to_vec2(tex_size);                            // :31
          Lane L;
          march(c, pixel_pos, screen_uv, L);
          lanes[ly * TILE_SIZE + lx] = L;
        }
      for (int ly = 0; ly < TILE_SIZE; ly++)
        for (int lx = 0; lx < TILE_SIZE; lx++) {
          ivec2 pixel_pos(tx * TILE_SIZE + lx, ty * TILE_SIZE + ly);
          const Lane& L = lanes[ly * TILE_SIZE + lx];
          if (L.sky) { imageStore(out, 0, pixel_pos, vec4(0.f, 0.f, 0.f, 1)); continue; }     // :240-242
          vec3 W0 = -normalize(L.camera_pos);                                                   // :308-333
          float weight = 0.f;
          float roughness = texture(c.material, L.screen_pos.xy()).y;
          vec3 accum(0, 0, 0);
          for (int x = lx - 1; x <= lx + 1; x++)
            for (int y = ly - 1; y <= ly + 1; y++)
              if (x >= 0 && x < TILE_SIZE && y >= 0 && y < TILE_SIZE) {
                const Lane& S = lanes[y * TILE_SIZE + x];
                vec3 ray_hit = S.hit_pos;
                if (ray_hit.z >= 0) {
                  vec3 camera_hit_pos = reconstruct_view_vec(ray_hit.xy(), ray_hit.z, c.cam);
                  vec3 Lv = normalize(camera_hit_pos - L.camera_pos);
                  vec3 H = normalize(W0 + Lv);
                  float w = DistributionGGX(L.camera_normal, H, roughness) * max(dot(L.camera_normal, Lv), 0.0f);
                  weight += w;
                  accum += S.hit_color * w;
                }
              }
          vec3 rgb = (weight > 0) ? (accum / weight) : vec3(0, 0, 0);
          float a = L.ao * (2 / 1.0f);                                                          // :338
          rgb /= 1.0f;
          imageStore(out, 0, pixel_pos, vec4(rgb, a));
        }
    }
  }
  return 0;
}

// filter.comp:13-39.  Push constants {znear, zfar}.  The sky store at :19-21 does not return: the 4x4 mean overwrites it.
extern "C" int orc_screen_trace_filter(float znear, float zfar, const OrcImage* raw_, const OrcImage* depth_, const OrcImage* out_, int y0, int y1) {
  const Image& out = *(const Image*)out_;
  View raw(*(const Image*)raw_), depth(*(const Image*)depth_, 0, 1);
  const int cw = (int)out.width / 8 * 8, ch = (int)out.height / 4 * 4;                         // cmd.dispatch(extent.width/8, extent.height/4, 1)
  if (y0 < 0) y0 = 0;
  if (y1 <= 0 || y1 > ch) y1 = ch;
#pragma omp parallel for schedule(dynamic, 8)
  for (int py = y0; py < y1; py++)
    for (int px = 0; px < cw; px++) {
      ivec2 pixel_pos(px, py);
      float pixel_depth = texelFetch(depth, pixel_pos, 0).x;
      float linear_depth = linearize_depth2(pixel_depth, znear, zfar);
      float weight_sum = 0;
      vec4 sum(0, 0, 0, 0);
      for (int x = 0; x < 4; x++)
        for (int y = 0; y < 4; y++) {
          ivec2 sample_pos = pixel_pos + ivec2(x - 2, y - 2);
          float sampled_depth = linearize_depth2(texelFetch(depth, sample_pos, 0).x, znear, zfar);
          float weight = max(0.0f, 1.0f - abs(sampled_depth - linear_depth) / (linear_depth * 0.1f));
          weight_sum += weight;
          sum = sum + weight * texelFetch(raw, sample_pos, 0);
        }
      sum = sum / weight_sum;
      imageStore(out, 0, pixel_pos, sum);
    }
  return 0;
}

// accumulate.comp:20-40.  Push constants {fovy, aspect, znear, zfar}; accum is read-modify-write.
extern "C" int orc_screen_trace_accumulate(const float* fovy_aspect_znear_zfar, const OrcImage* depth_, const OrcImage* prev_depth_,
                                           const OrcImage* filtered_, const OrcImage* accum_, int y0, int y1) {
  const Image& accum = *(const Image*)accum_;
  View depth(*(const Image*)depth_, 0, 1), prev_depth(*(const Image*)prev_depth_, 0, 1), filtered(*(const Image*)filtered_);
  const Cam cam(fovy_aspect_znear_zfar[0], fovy_aspect_znear_zfar[1], fovy_aspect_znear_zfar[2], fovy_aspect_znear_zfar[3]);
  const float REPROJECT_BIAS = 1e-6f, REPROJECT_COEF = 0.05f;
  const int cw = (int)accum.width / 8 * 8, ch = (int)accum.height / 4 * 4;
  const ivec2 tex_size(cw, ch);
  if (y0 < 0) y0 = 0;
  if (y1 <= 0 || y1 > ch) y1 = ch;
#pragma omp parallel for schedule(dynamic, 8)
  for (int py = y0; py < y1; py++)
    for (int px = 0; px < cw; px++) {
      ivec2 pixel_pos(px, py);
      vec2 screen_uv = to_vec2(pixel_pos) / to_vec2(tex_size);
      vec4 new_sum = texelFetch(filtered, pixel_pos, 0);
      float current_depth = texelFetch(depth, pixel_pos, 0).x;
      vec3 cur_view = reconstruct_view_vec(screen_uv, current_depth, cam);
      vec4 out_sum = new_sum;
      float sampled_depth = texelFetch(prev_depth, pixel_pos, 0).x;
      vec4 sampled_sum = imageLoad(accum, 0, pixel_pos);
      float sampled_z = linearize_depth2(sampled_depth, cam.znear, cam.zfar);
      float delta = abs(sampled_z - cur_view.z);
      if (delta < REPROJECT_BIAS && sampled_depth < 1.0f) out_sum = mix(sampled_sum, new_sum, REPROJECT_COEF);
      imageStore(accum, 0, pixel_pos, out_sum);
    }
  return 0;
}
