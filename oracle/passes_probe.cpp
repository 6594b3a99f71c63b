// ORACLE — TEST INFRASTRUCTURE ONLY (see oracle.h header).  PARITY UNPINNED.
//
// The probe_renderer compute path:
//   render_octahedral   reference src/probe_renderer.cpp:170-204, shaders/cube2oct/shader.comp:14-33, include/octahedral.glsl
//   probe_downsample    reference src/probe_renderer.cpp:206-236, shaders/probe_downsample/shader.frag:8-23
//   ProbeTracePass::run reference src/probe_renderer.cpp:316-385, shaders/trace_probe/shader.comp
//
// Cube-map sampling is Vulkan fixed function (face selection table + bilinear filtering).  Interpretation pinned here: the
// major axis picks the face (ties: z over y over x), (s, t) = 0.5 * (sc, tc) / |rc| + 0.5, and the bilinear footprint is clamped
// to the face (no blending across face edges; a real device blends the outermost half texel with the neighbouring face).
#include <cmath>
#include "oracle.h"
#include "shader_include.hpp"

using namespace orc;

namespace {

// ---- octahedral.glsl ---------------------------------------------------------------------------------------------
float o_sign_nz(float k) { return (k >= 0.0f) ? 1.0f : -1.0f; }
float sign_g(float x) { return x > 0.0f ? 1.0f : (x < 0.0f ? -1.0f : 0.0f); }

vec3 oct_unfold(vec2 uv) {                                                                       // shared head of oct_decode / oct_center (:38-45)
  uv = 2.f * (uv - vec2(0.5f, 0.5f));
  vec3 v = vec3(uv.x, uv.y, 1.0f - abs(uv.x) - abs(uv.y));
  if (v.z < 0.0f) {
    float x = (1.0f - abs(v.y)) * o_sign_nz(v.x), y = (1.0f - abs(v.x)) * o_sign_nz(v.y);
    v.x = x; v.y = y;
  }
  return v;
}
vec3 oct_decode(vec2 uv) { return normalize(oct_unfold(uv)); }                                    // :37-46
vec3 oct_center(vec2 uv) { vec3 v = oct_unfold(uv); return normalize(vec3(sign_g(v.x), sign_g(v.y), sign_g(v.z))); }   // :48-57
vec2 oct_encode(vec3 v) {                                                                         // :59-66
  float l1norm = abs(v.x) + abs(v.y) + abs(v.z);
  vec2 result = v.xy() * (1.0f / l1norm);
  if (v.z < 0.0f) result = (1.0f - abs(vec2(result.y, result.x))) * vec2(o_sign_nz(result.x), o_sign_nz(result.y));
  return 0.5f * result + vec2(0.5f, 0.5f);
}
float encode_oct_depth(float z, float n, float f) { return f / (f - n) + f * n / ((-z) * (f - n)); }   // :73-75
float decode_oct_depth(float d, float n, float f) { return -n * f / (d * (f - n) - f); }               // :77-79

const float ZNEAR = 0.05f, ZFAR = 80.f;

Image layer_image(const Image& im, uint64_t layer_stride, uint32_t layer) {
  Image l = im;
  l.base = (uint8_t*)im.base + layer_stride * layer;
  return l;
}

// texture(samplerCube, dir): see the header for the pinned interpretation.
vec4 texture_cube(const Image& face0, uint64_t face_stride, vec3 r) {
  float ax = abs(r.x), ay = abs(r.y), az = abs(r.z);
  int face; float sc, tc, rc;
  if (az >= ax && az >= ay) { face = r.z < 0 ? 5 : 4; sc = r.z < 0 ? -r.x : r.x; tc = -r.y; rc = az; }
  else if (ay >= ax)        { face = r.y < 0 ? 3 : 2; sc = r.x; tc = r.y < 0 ? -r.z : r.z; rc = ay; }
  else                      { face = r.x < 0 ? 1 : 0; sc = r.x < 0 ? r.z : -r.z; tc = -r.y; rc = ax; }
  vec2 st(0.5f * (sc / rc) + 0.5f, 0.5f * (tc / rc) + 0.5f);
  Image f = layer_image(face0, face_stride, (uint32_t)face);
  return texture(View(f), st);
}

}  // namespace

// oct_color / oct_depth: the target array layer as an image of its own (R8G8B8A8_UNORM; R16_UNORM, mip 0 written).
extern "C" int orc_cube2oct(const OrcImage* cube_color_, uint64_t color_face_stride, const OrcImage* cube_distance_, uint64_t distance_face_stride,
                            const OrcImage* oct_color_, const OrcImage* oct_depth_) {
  const Image &cc = *(const Image*)cube_color_, &cd = *(const Image*)cube_distance_, &oc = *(const Image*)oct_color_, &od = *(const Image*)oct_depth_;
  const int cw = (int)oc.width / 8 * 8, ch = (int)oc.height / 4 * 4;                               // probe_renderer.cpp:201
  const ivec2 tex_size(cw, ch);
#pragma omp parallel for schedule(dynamic, 8)
  for (int py = 0; py < ch; py++)
    for (int px = 0; px < cw; px++) {
      ivec2 pixel_pos(px, py);
      vec2 screen_uv = to_vec2(pixel_pos) / to_vec2(tex_size);
      vec3 cube_direction = oct_decode(screen_uv);
      vec4 sampled_color = texture_cube(cc, color_face_stride, cube_direction);
      float sampled_distance = texture_cube(cd, distance_face_stride, cube_direction).x;
      vec3 view_dir = cube_direction * sampled_distance;
      vec3 front = oct_center(screen_uv);
      float depth = encode_oct_depth(clamp(dot(view_dir, front), ZNEAR, ZFAR), ZNEAR, ZFAR);
      imageStore(oc, 0, pixel_pos, sampled_color);
      imageStore(od, 0, pixel_pos, vec4(depth, 0, 0, 0));
    }
  return 0;
}

// Every mip i >= 1 of one layer from mip i-1.  The clamp `min(pixel + off, size)` (:17-20) admits index == size, which
// texelFetch answers with 0 — as written.
extern "C" int orc_probe_downsample(const OrcImage* probe_depth_layer_) {
  const Image& im = *(const Image*)probe_depth_layer_;
  for (uint32_t i = 1; i < im.mip_count; i++) {
    View src(im, (int)i - 1, 1);
    const int w = (int)im.width / (1 << i), h = (int)im.height / (1 << i);                         // :226
    ivec2 size = textureSize(src, 0);
    for (int py = 0; py < h; py++)
      for (int px = 0; px < w; px++) {
        float min_depth = 1000.f;
        ivec2 pixel(2 * px, 2 * py);
        const int offs[4][2] = {{0, 0}, {0, 1}, {1, 0}, {1, 1}};
        for (auto& o : offs) {
          ivec2 q(min(pixel.x + o[0], size.x), min(pixel.y + o[1], size.y));
          min_depth = min(texelFetch(src, q, 0).x, min_depth);
        }
        imageStore(im, (int)i, ivec2(px, py), vec4(min_depth, 0, 0, 0));
      }
  }
  return 0;
}

// ---- trace_probe/shader.comp --------------------------------------------------------------------------------------
namespace {

enum { TRACE_RESULT_MISS = 0, TRACE_RESULT_HIT = 1, TRACE_RESULT_UNKNOWN = 2 };
const uint32_t TRACE_STEPS = 25;

struct ProbeCtx {
  Image probe_color, probe_depth;            // layer 0; layers lie *_stride bytes apart
  uint64_t color_stride, depth_stride;
  vec3 probe_min, probe_max;
  uint32_t grid_size;
};

float dist_squared(vec3 v0, vec3 v1) { vec3 t = v0 - v1; return dot(t, t); }                      // :98-101
void swap_min(float& a, float& b) { float temp = min(a, b); b = max(a, b); a = temp; }            // :104-108

void compute_trace_segments(vec3 origin, vec3 dir_frac, float tmin, float tmax, float bounds[5]) {   // :110-124
  bounds[0] = tmin;
  vec3 t = origin * (-dir_frac);
  swap_min(t.x, t.y);
  swap_min(t.y, t.z);
  swap_min(t.x, t.y);
  bounds[1] = clamp(t.x, tmin, tmax);
  bounds[2] = clamp(t.y, tmin, tmax);
  bounds[3] = clamp(t.z, tmin, tmax);
  bounds[4] = tmax;
}

// :191-218 — this shader's own copy: t_min is additionally capped at 1 (the end of the segment)
bool advance_ray_p(vec3 origin, vec3 direction, vec3 inv_direction, vec2 current_mip_position, vec2 current_mip_resolution_inv, vec2 floor_offset,
                   vec2 uv_offset, float surface_z, vec3& position, float& current_t) {
  vec2 xy_plane = floor(current_mip_position) + floor_offset;
  xy_plane = xy_plane * current_mip_resolution_inv + uv_offset;
  vec3 boundary_planes = vec3(xy_plane, surface_z);
  vec3 t = (boundary_planes - origin) * inv_direction;
  t.z = direction.z > 0 ? t.z : MAX_T_FLOAT;
  float t_min = min(min(min(t.x, t.y), t.z), 1.0f);
  bool above_surface = surface_z > position.z;
  bool skipped_tile = t_min != t.z && above_surface;
  current_t = above_surface ? t_min : current_t;
  position = origin + current_t * direction;
  return skipped_tile;
}

// :224-275
vec3 hierarchical_raymarch_p(const ProbeCtx& c, int probe_index, vec3 origin, vec3 direction, int most_detailed_mip, uint32_t max_traversal_intersections,
                             bool& valid_hit) {
  Image layer = layer_image(c.probe_depth, c.depth_stride, (uint32_t)probe_index);
  View depth_tex(layer);
  const vec3 inv_direction = vec3(direction.x != 0 ? 1.0f / direction.x : MAX_T_FLOAT, direction.y != 0 ? 1.0f / direction.y : MAX_T_FLOAT,
                                  direction.z != 0 ? 1.0f / direction.z : MAX_T_FLOAT);
  int current_mip = 0;
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
  uint32_t i = 0;
  while (i < max_traversal_intersections && current_mip >= most_detailed_mip) {
    vec2 current_mip_position = current_mip_resolution * position.xy();
    float surface_z = texelFetch(depth_tex, to_ivec2(current_mip_position), current_mip).x;
    bool skipped_tile = advance_ray_p(origin, direction, inv_direction, current_mip_position, current_mip_resolution_inv, floor_offset, uv_offset, surface_z,
                                      position, current_t);
    current_mip += skipped_tile ? 1 : -1;
    current_mip_resolution *= skipped_tile ? 0.5f : 2.0f;
    current_mip_resolution_inv *= skipped_tile ? 2.0f : 0.5f;
    ++i;
  }
  valid_hit = (i <= max_traversal_intersections);
  return position;
}

// :277-331
int trace_segment_hi(const ProbeCtx& c, vec3 ray_origin, vec3 ray_dir, float t0, float t1, float& tmin, int probe_index, vec2& stop_tc) {
  const float RAY_EPS = 0.001f;
  vec3 probe_start = ray_origin + ray_dir * (t0 + RAY_EPS);
  vec3 probe_end = ray_origin + ray_dir * (t1 - RAY_EPS);
  if (dist_squared(probe_start, probe_end) < 0.001f) probe_start = ray_dir;
  vec2 start_oct = oct_encode(normalize(probe_start));
  vec2 end_oct = oct_encode(normalize(probe_end));
  vec3 front = oct_center(0.5f * (start_oct + end_oct));
  const float RAY_START_BIAS = 0.0005f;
  float start_depth = encode_oct_depth(dot(probe_start, front), ZNEAR, ZFAR) - RAY_START_BIAS;
  float end_depth = encode_oct_depth(dot(probe_end, front), ZNEAR, ZFAR);
  vec3 p_start = vec3(start_oct, start_depth);
  vec3 p_end = vec3(end_oct, end_depth);
  bool valid = false;
  vec3 p_stop = hierarchical_raymarch_p(c, probe_index, p_start, p_end - p_start, 0, TRACE_STEPS, valid);
  const float HIT_BIAS = 0.0005f;
  if (valid) {
    stop_tc = p_stop.xy();
    vec3 stop_dir = oct_decode(stop_tc);
    float coef = decode_oct_depth(p_stop.z, ZNEAR, ZFAR) / dot(stop_dir, front);
    vec3 ray_stop = stop_dir * coef;
    tmin = length(ray_stop - ray_origin);
    Image layer = layer_image(c.probe_depth, c.depth_stride, (uint32_t)probe_index);
    float sampled_depth = textureLod(View(layer), p_stop.xy(), 0).x;
    if (p_stop.z > 1) return TRACE_RESULT_MISS;
    if (p_stop.z > sampled_depth + HIT_BIAS) return TRACE_RESULT_UNKNOWN;
    if (p_stop.z > sampled_depth - HIT_BIAS) return TRACE_RESULT_HIT;
  }
  return TRACE_RESULT_MISS;
}

int get_start_probe_index(const ProbeCtx& c, vec3 world_pos) {                                    // :333-339
  vec3 probe_step = (c.probe_max - c.probe_min) / float(c.grid_size - 1);
  vec3 coord = clamp((world_pos - c.probe_min) / probe_step, vec3(0, 0, 0), vec3(float(c.grid_size - 2), 2, float(c.grid_size - 2)));
  int x = int(floor(coord.x));
  int y = int(floor(coord.z));
  return y * int(c.grid_size) + x;
}

vec3 get_probe_position(const ProbeCtx& c, int index) {                                           // :341-346
  float y = float(uint32_t(index) / c.grid_size);
  float x = float(uint32_t(index) % c.grid_size);
  vec3 probe_step = (c.probe_max - c.probe_min) / float(c.grid_size - 1);
  return c.probe_min + vec3(x, 0, y) * probe_step;
}

int next_probe(const ProbeCtx& c, int start_probe, int i) {                                       // :348-354
  int x = start_probe % int(c.grid_size);
  int y = start_probe / int(c.grid_size);
  int dy = (i >> 1) & 1;
  int dx = i & 1;
  return (y + dy) * int(c.grid_size) + x + dx;
}

// :386-412 (the overload trace() calls)
int trace_probe(const ProbeCtx& c, vec3 ray_origin, vec3 ray_dir, int probe_index, float& tmin, float& tmax, vec2& hitexc) {
  const float degenerateEpsilon = 0.002f;
  vec3 probeOrigin = get_probe_position(c, probe_index);
  vec3 probeRayOrigin = ray_origin - probeOrigin;
  vec3 probeRayDir = normalize(ray_dir);
  float segments[5];
  compute_trace_segments(probeRayOrigin, 1.0f / probeRayDir, tmin, tmax, segments);
  for (int i = 0; i < 4; i++)
    if (abs(segments[i + 1] - segments[i]) >= degenerateEpsilon) {
      int result = trace_segment_hi(c, probeRayOrigin, probeRayDir, segments[i], segments[i + 1], tmin, probe_index, hitexc);
      if (result != TRACE_RESULT_MISS) return result;
    }
  return TRACE_RESULT_MISS;
}

// :414-444
int trace(const ProbeCtx& c, vec3 ray_origin, vec3 ray_dir, float& tmax, vec2& hittexc, int& hit_probe) {
  hit_probe = -1;
  int base_probe = get_start_probe_index(c, ray_origin);
  int i = 0;
  int probes_left = 4;
  float tmin = 0.f;
  int result = TRACE_RESULT_UNKNOWN;
  while (probes_left > 0) {
    int probe_id = next_probe(c, base_probe, i);
    result = trace_probe(c, ray_origin, ray_dir, probe_id, tmin, tmax, hittexc);
    if (result == TRACE_RESULT_UNKNOWN) {
      i = (i + 3) & 3;
      probes_left--;
    } else {
      if (result == TRACE_RESULT_HIT) hit_probe = probe_id;
      break;
    }
  }
  return result;
}

}  // namespace

// main(), :48-86.  out = R8G8B8A8_UNORM; dispatch floor(W/8) x floor(H/4) groups (probe_renderer.cpp:381).
extern "C" int orc_trace_probe(const OrcProbeTraceParams* p, const OrcImage* depth_, const OrcImage* normal_, const OrcImage* probe_color_,
                               uint64_t color_layer_stride, const OrcImage* probe_depth_, uint64_t depth_layer_stride, const OrcImage* out_, int y0, int y1) {
  const Image& out = *(const Image*)out_;
  View DEPTH_TEX(*(const Image*)depth_, 0, 1), NORMAL_TEX(*(const Image*)normal_);
  ProbeCtx c{*(const Image*)probe_color_, *(const Image*)probe_depth_, color_layer_stride, depth_layer_stride,
             vec3(p->probe_min[0], p->probe_min[1], p->probe_min[2]), vec3(p->probe_max[0], p->probe_max[1], p->probe_max[2]), p->grid_size};
  const mat4 inverse_view(p->inverse_view);
  const Cam cam(p->fovy, p->aspect, p->znear, p->zfar);
  const int cw = (int)out.width / 8 * 8, ch = (int)out.height / 4 * 4;
  const ivec2 tex_size(cw, ch);
  if (y0 < 0) y0 = 0;
  if (y1 <= 0 || y1 > ch) y1 = ch;
#pragma omp parallel for schedule(dynamic, 4)
  for (int py = y0; py < y1; py++)
    for (int px = 0; px < cw; px++) {
      ivec2 pixel_pos(px, py);
      vec2 screen_uv = to_vec2(pixel_pos) / to_vec2(tex_size);
      float pixel_depth = texture(DEPTH_TEX, screen_uv).x;
      if (pixel_depth >= 1.f) { imageStore(out, 0, pixel_pos, vec4(0, 0, 0, 0)); continue; }
      vec3 view_vec = reconstruct_view_vec(screen_uv, pixel_depth, cam);
      vec3 N = sample_gbuffer_normal(NORMAL_TEX, screen_uv);
      vec3 world_pos = (inverse_view * vec4(view_vec, 1)).xyz();
      world_pos += 1e-6f * N;
      vec3 camera_pos = (inverse_view * vec4(0, 0, 0, 1)).xyz();
      vec3 V = normalize(world_pos - camera_pos);
      world_pos -= 1e-6f * V;
      vec3 R = reflect(V, N);
      vec4 reflection(0, 0, 0, 0);
      float tmax = 30.f;
      vec2 hit_pos;
      int probe_index;
      int res = trace(c, world_pos, R, tmax, hit_pos, probe_index);
      if (res == TRACE_RESULT_HIT) {
        Image layer = layer_image(c.probe_color, c.color_stride, (uint32_t)probe_index);
        reflection = texture(View(layer), hit_pos);
      }
      imageStore(out, 0, pixel_pos, reflection);
    }
  return 0;
}
