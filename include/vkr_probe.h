/* vkr_probe.h — C ABI of the probe_renderer compute path: cube map -> octahedral probe, the probe's min-depth
 * pyramid, and the reflection trace through a grid of octahedral probes.
 *
 * Replaces the task bodies of reference src/probe_renderer.cpp: "Cubemap2Octahedral" (:170-204, shaders/cube2oct/shader.comp),
 * "DownsampleProbe" (:206-236, shaders/probe_downsample/shader.frag) and "TraceProbe" (ProbeTracePass::run :316-385,
 * shaders/trace_probe/shader.comp).  The cube faces themselves are rasterised by the reference (render_cubemap :78-168): they
 * are INPUTS here, like the G-buffer.
 *
 * Array images are a VkrImage describing layer (or face) 0 plus the byte distance between layers; a cube map is six faces in
 * Vulkan order +x -x +y -y +z -z.  Cube sampling = face selection by major axis (ties z over y over x) + bilinear filtering
 * clamped to the face.  Same conventions as vkr_passes.h: borrowed descriptors, asynchronous on `stream`, 0 / negative VkrStatus.
 */
#ifndef VKR_PROBE_H_INCLUDED
#define VKR_PROBE_H_INCLUDED
#include "vkr_passes.h"

#ifdef __cplusplus
extern "C" {
#endif

/* Constants UBO of ProbeTracePass::run (reference src/probe_renderer.cpp:340-360). */
typedef struct VkrProbeTraceParams {
  float inverse_view[16];     /* camera -> world */
  float probe_min[4];         /* xyz = first probe of the grid (w unused) */
  float probe_max[4];         /* xyz = last probe */
  uint32_t grid_size;         /* grid_size x grid_size probes in the xz plane, layer = z_index * grid_size + x_index */
  float fovy, aspect, znear, zfar;
} VkrProbeTraceParams;

/* cube_color: R8G8B8A8_SRGB faces, cube_distance: R16_SFLOAT faces (radial distance, cubemap_probe/shader.frag:32);
 * oct_color: R8G8B8A8_UNORM, oct_depth: R16_UNORM (mip 0 is written).  Converts probe_count probes in one launch: probe i reads
 * the cube set at base + i * 6 * face_stride and writes layer first_layer + i.  The reference converts one per task (:192). */
int vkr_cube2oct(const VkrImage* cube_color, uint64_t color_face_stride, const VkrImage* cube_distance, uint64_t distance_face_stride,
                 const VkrImage* oct_color, uint64_t oct_color_layer_stride, const VkrImage* oct_depth, uint64_t oct_depth_layer_stride,
                 uint32_t first_layer, uint32_t probe_count, void* stream);
/* Mips 1..L-1 of layers [first_layer, first_layer + layer_count) of an R16_UNORM probe depth array, each from the previous
 * (2x2 min; mip i is (W >> i) x (H >> i), :226). */
int vkr_probe_downsample(const VkrImage* probe_depth, uint64_t layer_stride, uint32_t first_layer, uint32_t layer_count, void* stream);
/* gbuffer depth (mip 0) + world-space oct normals -> out R8G8B8A8_UNORM reflection colour (0 where nothing is hit or the pixel
 * is sky).  Only floor(W/8)*8 x floor(H/4)*4 pixels are written (:381).  ext (may be NULL): row window. */
int vkr_trace_probe(const VkrProbeTraceParams* p, const VkrImage* depth, const VkrImage* normal, const VkrImage* probe_color,
                    uint64_t color_layer_stride, const VkrImage* probe_depth, uint64_t depth_layer_stride, const VkrImage* out,
                    const VkrLaunchExt* ext, void* stream);

#ifdef __cplusplus
}
#endif
#endif
