/* vkr_synth.h — deterministic synthetic G-buffer producer (CUDA ray-caster + host twin).
 *
 * Stands in for the rasterised G-buffer pass (task "GbufferPass", reference
 * src/scene_renderer.cpp:140-220, shaders/gbuf/opaque_taa.{vert,frag}), which needs a
 * graphics API and a glTF scene and is therefore out of scope.  It
 * writes exactly the images `struct Gbuffer` holds (reference src/scene_renderer.hpp:11-24)
 * in the reference's formats and conventions:
 *   depth    D24 in the low 24 bits, d = f/(f-n) + f*n/(z*(f-n))   (gbuffer_encode.glsl:71-73)
 *   normal   R16G16_UNORM octahedral world normal                  (gbuffer_encode.glsl:17-27)
 *   albedo / material  R8G8B8A8_SRGB  (material .g roughness, .b metallic)
 *   velocity R16G16_SFLOAT, 0.5*(prev_ndc.xy - ndc.xy)             (gbuf/opaque_taa.frag:45)
 * The scene (ground plane, back wall, spheres and boxes) and the camera path are pure
 * functions of (seed, frame) built on the host from integer hashes; the per-pixel code is
 * one __host__ __device__ function, so the CUDA kernel and the host twin produce the same
 * images (used by tests to feed the CPU oracle and the CUDA passes identical inputs).
 */
#ifndef VKR_SYNTH_H_INCLUDED
#define VKR_SYNTH_H_INCLUDED
#include "vkr_passes.h"

#ifdef __cplusplus
extern "C" {
#endif

#define VKR_SYNTH_MAX_OBJECTS 64

typedef struct VkrSynthObject {
  float center[3];
  float half_extent[3];   /* sphere: radius in [0] */
  uint32_t kind;          /* 0 = sphere, 1 = axis-aligned box */
  uint32_t albedo_srgb8;  /* packed R8G8B8A8 code written to gbuffer.albedo */
  uint32_t material_srgb8;/* packed code written to gbuffer.material */
} VkrSynthObject;

typedef struct VkrSynthScene {
  uint32_t object_count;
  VkrSynthObject objects[VKR_SYNTH_MAX_OBJECTS];
  uint32_t ground_albedo[2], ground_material;  /* checkerboard */
  uint32_t wall_albedo, wall_material;
  uint32_t sky_albedo;
  float wall_z, wall_height, wall_half_width;
} VkrSynthScene;

/* Mirrors DrawTAAParams (reference src/scene_renderer.hpp:26-33) as built per frame in
 * src/main.cpp:329-336.  Column-major mat4s. */
typedef struct VkrDrawTAAParams {
  float mvp[16];
  float prev_mvp[16];
  float camera[16];
  float prev_camera[16];
  float jitter[4];
  float fovy_aspect_znear_zfar[4];
} VkrDrawTAAParams;

/* Host: scene from a seed (no device work). */
void vkr_synth_scene(uint32_t seed, uint32_t object_count, VkrSynthScene* out);
/* Host: camera of `frame` on the synthetic path (translation 0.02/frame + 0.1 deg/frame yaw),
 * fov 60 deg, znear 0.05, zfar 80 (reference src/main.cpp:294,:335), aspect = width/height,
 * 4-sample TAA jitter (src/main.cpp:93-108) when use_jitter != 0.  prev_* = camera of
 * frame-1 (frame 0: prev = current, src/main.cpp:296-298).  `view_seed` picks one of many start poses. */
void vkr_synth_camera(uint32_t view_seed, uint32_t frame, uint32_t width, uint32_t height, int use_jitter, VkrDrawTAAParams* out);
/* Host helpers the pass classes need (glm::inverse / transpose(inverse) of a rigid view matrix). */
void vkr_mat4_inverse(const float* m, float* out);
void vkr_mat4_transpose(const float* m, float* out);

/* Device: ray-cast one frame into device images (asynchronous on stream). depth: mip 0 written. */
int vkr_synth_gbuffer(const VkrSynthScene* scene, const VkrDrawTAAParams* params, const VkrImage* albedo,
                      const VkrImage* normal, const VkrImage* material, const VkrImage* velocity,
                      const VkrImage* depth, void* stream);
/* Host twin: same images in host memory, rows split over `threads` std::threads (0 = all cores). */
int vkr_synth_gbuffer_host(const VkrSynthScene* scene, const VkrDrawTAAParams* params, const VkrImage* albedo,
                           const VkrImage* normal, const VkrImage* material, const VkrImage* velocity,
                           const VkrImage* depth, int threads);

/* Device: the six cube faces of a probe at `position3` (stand-in for ProbeRenderer::render_cubemap, reference
 * src/probe_renderer.cpp:78-168, shaders/cubemap_probe/shader.frag): cube_color R8G8B8A8_SRGB = albedo, cube_distance
 * R16_SFLOAT = radial distance, faces in Vulkan order +x -x +y -y +z -z at base + f * face_stride. */
int vkr_synth_cubemap(const VkrSynthScene* scene, const float* position3, const VkrImage* cube_color, uint64_t color_face_stride,
                      const VkrImage* cube_distance, uint64_t distance_face_stride, void* stream);

#ifdef __cplusplus
}
#endif
#endif
