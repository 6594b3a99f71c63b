// march.cuh — the Hi-Z cell march shared by the stochastic SSR trace and the legacy SSR pass
// (reference src/shaders/include/screen_trace.glsl:4-49).  Everything here decides texel indices, so it is
// evaluated unfused with IEEE divisions (common.cuh precision policy).
#pragma once
#include "common.cuh"

namespace vkr {

struct PyramidView {   // a mip-range view of the depth image: view lod m = image mip base + m
  Mip mip[VKR_MAX_MIPS];
  int count;
};

#define MAX_T_FLOAT 3.402823466e+38f

__device__ __forceinline__ float fetch_pyramid(const PyramidView& P, unsigned long long layer_off, int x, int y, int lod) {
  if ((unsigned)lod >= (unsigned)P.count) return 0.0f;
  const Mip& m = P.mip[lod];
  if ((unsigned)x >= (unsigned)m.w || (unsigned)y >= (unsigned)m.h) return 0.0f;      // texelFetch out of range = 0 (App. A.3)
  return d24_to_float(__ldg((const uint32_t*)(m.p + layer_off + (size_t)y * m.pitch) + x));
}

// screen_trace.glsl:17-45
__device__ __forceinline__ bool advance_ray(f3 origin, f3 direction, f3 inv_direction, float mpx, float mpy, float rinvx, float rinvy,
                                            float fox, float foy, float uox, float uoy, float surface_z, f3& position, float& current_t) {
  float xp = FA(FM(FA(floorf(mpx), fox), rinvx), uox);
  float yp = FA(FM(FA(floorf(mpy), foy), rinvy), uoy);
  float tx = FM(FS(xp, origin.x), inv_direction.x);
  float ty = FM(FS(yp, origin.y), inv_direction.y);
  float tz = FM(FS(surface_z, origin.z), inv_direction.z);
  tz = direction.z > 0 ? tz : MAX_T_FLOAT;
  float t_min = gmin(gmin(tx, ty), tz);
  bool above_surface = surface_z > position.z;
  bool skipped_tile = t_min != tz && above_surface;
  current_t = above_surface ? t_min : current_t;
  position = origin + current_t * direction;
  return skipped_tile;
}

}  // namespace vkr
