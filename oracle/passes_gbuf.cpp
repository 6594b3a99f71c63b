// ORACLE — TEST INFRASTRUCTURE ONLY (see oracle.h header).
//
// The conventions a G-buffer producer must honour for the image-space chain, restated from the reference's rasterised
// G-buffer pass.  Not a rasteriser: the per-vertex / per-fragment arithmetic of the three places that define how jitter and
// motion vectors are written, evaluated for one world-space point.  tests/test_producer_conventions*.py hold the CUDA
// producer (vk-renderer_b200/csrc/synth.cu) and its host twin to these.
//   reference src/shaders/gbuf/opaque_taa.vert:37-43   out_vector = view_projection * model * pos;
//                                                      gl_Position = out_vector + out_vector.w * vec4(jitter.xy, 0, 0);
//                                                      pos_after = out_vector; pos_before = prev_view_projection * model * pos
//   reference src/shaders/gbuf/opaque_taa.frag:45      velocity_vector = 0.5 * (pos_before.xy/pos_before.w - pos_after.xy/pos_after.w)
//   reference src/main.cpp:93-108                      next_taa_offset: offsets[index], offset = 2*o - 1, times 1/resolution, index = (index+1) % 4
#include "oracle.h"
#include "shader_include.hpp"

using namespace orc;

extern "C" {

// main.cpp:93-108.  `index` is the value of the function's static counter at the call (0 for the first frame).
void orc_taa_offset(uint32_t index, uint32_t w, uint32_t h, float* out_jitter4) {
  static const float offsets[4][2] = {{0.25f, 0.25f}, {0.75f, 0.75f}, {0.75f, 0.25f}, {0.25f, 0.75f}};
  const vec2 inv_resolution((float)(1.0 / w), (float)(1.0 / h));           // glm::vec2{1.0/w, 1.0/h}: double quotient narrowed
  const vec2 o(offsets[index % 4][0], offsets[index % 4][1]);
  const vec2 offset = 2.0f * o - vec2(1.0f, 1.0f);
  const vec2 r = offset * inv_resolution;
  out_jitter4[0] = r.x; out_jitter4[1] = r.y; out_jitter4[2] = 0.f; out_jitter4[3] = 0.f;
}

// opaque_taa.vert:37-43 + opaque_taa.frag:45 for one point already in world space (model = identity).
// out_position4 = gl_Position (jittered clip coordinates), out_velocity2 = velocity_vector.
void orc_gbuf_point(const float* view_projection16, const float* prev_view_projection16, const float* jitter4,
                    const float* world_pos3, float* out_position4, float* out_velocity2) {
  const mat4 view_projection(view_projection16), prev_view_projection(prev_view_projection16);
  const vec4 in_pos(world_pos3[0], world_pos3[1], world_pos3[2], 1.0f);
  const vec4 out_vector = view_projection * in_pos;
  const vec4 gl_Position = out_vector + vec4(jitter4[0], jitter4[1], 0.0f, 0.0f) * out_vector.w;
  const vec4 pos_after = out_vector;
  const vec4 pos_before = prev_view_projection * in_pos;
  const vec2 velocity_vector = 0.5f * (vec2(pos_before.x, pos_before.y) / pos_before.w - vec2(pos_after.x, pos_after.y) / pos_after.w);
  out_position4[0] = gl_Position.x; out_position4[1] = gl_Position.y; out_position4[2] = gl_Position.z; out_position4[3] = gl_Position.w;
  out_velocity2[0] = velocity_vector.x; out_velocity2[1] = velocity_vector.y;
}

}  // extern "C"
