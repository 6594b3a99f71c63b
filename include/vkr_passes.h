/* vkr_passes.h — C ABI of the B200 (sm_100a) image-space passes.
 *
 * One entry point per rendergraph task of vk-renderer's image-space chain.  Each entry is
 * what the task's exec callback (the lambda handed to RenderGraph::add_task, reference
 * src/rendergraph/rendergraph.hpp:116-128) calls instead of recording
 * bind_pipeline / write_set / push_constants / dispatch|draw on a gpu::CmdContext.
 *
 * Conventions:
 *   - images are headless linear device buffers described by VkrImage; the callee BORROWS
 *     them (never frees, never retains past the call's stream work);
 *   - params structs mirror the UBO + push-constant layout of the reference task
 *     (column-major mat4 as float[16], exactly what glm::mat4 holds);
 *   - every call is asynchronous on `stream` (a cudaStream_t passed as void*; NULL = the
 *     legacy default stream), holds no mutable global state and is re-entrant per stream;
 *   - returns 0 on success, negative VkrStatus on failure; no exceptions cross the ABI.
 *     The C++ host shim turns non-zero into std::runtime_error, the reference's error
 *     convention (src/gpu/common.hpp:9, src/downsample_pass.cpp:38,:49).
 *   - there is NO CPU fallback: without a CUDA device every compute entry returns
 *     VKR_ERR_CUDA.
 */
#ifndef VKR_PASSES_H_INCLUDED
#define VKR_PASSES_H_INCLUDED
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define VKR_MAX_MIPS 16

typedef enum VkrStatus {
  VKR_OK = 0,
  VKR_ERR_INVALID_ARGUMENT = -1, /* null pointer, wrong format, bad mip window */
  VKR_ERR_SIZE_MISMATCH = -2,    /* reference: "Output textures have different sizes" */
  VKR_ERR_CUDA = -3,             /* launch / runtime error; see vkr_last_error() */
  VKR_ERR_UNSUPPORTED = -4
} VkrStatus;

/* Values are VkFormat numbers so descriptors read like gpu::ImageInfo
 * (reference src/scene_renderer.cpp:15-27, src/gtao.cpp:26-39, src/advanced_ssr.cpp:62-92). */
typedef enum VkrFormat {
  VKR_FORMAT_R8_UNORM = 9,
  VKR_FORMAT_R8G8B8A8_UNORM = 37,
  VKR_FORMAT_R8G8B8A8_SRGB = 43,
  VKR_FORMAT_R16_UNORM = 70,
  VKR_FORMAT_R16_SFLOAT = 76,
  VKR_FORMAT_R16G16_UNORM = 77,
  VKR_FORMAT_R16G16_SFLOAT = 83,
  VKR_FORMAT_R16G16B16A16_UNORM = 91,
  VKR_FORMAT_R16G16B16A16_SFLOAT = 97,
  VKR_FORMAT_R32_SFLOAT = 100,
  VKR_FORMAT_R32G32B32A32_SFLOAT = 109,
  VKR_FORMAT_D24_UNORM_S8_UINT = 129 /* u32 texel, depth in the low 24 bits (src/main.cpp:139) */
} VkrFormat;

/* Replaces gpu::ImagePtr + gpu::ImageViewRange (reference src/gpu/resource_info.hpp:11-45).
 * One allocation holds the whole mip chain. */
typedef struct VkrImage {
  void* base;                       /* device pointer */
  uint32_t width, height;           /* mip 0 extent; mip i = max(1, dim >> i) */
  uint32_t format;                  /* VkrFormat */
  uint32_t mip_count;
  uint64_t mip_offset[VKR_MAX_MIPS]; /* bytes from base */
  uint32_t row_pitch[VKR_MAX_MIPS];  /* bytes; multiple of 16 */
} VkrImage;

/* SSSR_trace / SSSR_filter: UBO TraceParams (reference src/advanced_ssr.cpp:138-145) +
 * push constants {max_roughness} (:159-163) / {render_flags} (:328-338). */
#define VKR_SSR_NORMALIZE_REFLECTIONS 1u
#define VKR_SSR_ACCUMULATE_REFLECTIONS 2u
#define VKR_SSR_BILATERAL_FILTER 4u
typedef struct VkrTraceParams {
  float normal_mat[16];
  uint32_t frame_random;
  float fovy, aspect, znear, zfar;
  float max_roughness;
  uint32_t render_flags;
} VkrTraceParams;

/* SSSR_blur: UBO {inverse_camera, prev_inverse_camera, fovy_aspect_znear_zfar} +
 * push {max_roughness, accumulate, disable_blur} (reference src/advanced_ssr.cpp:388-401). */
typedef struct VkrBlurParams {
  float inverse_camera[16];
  float prev_inverse_camera[16];
  float fovy_aspect_znear_zfar[4];
  float max_roughness;
  uint32_t accumulate;
  uint32_t disable_blur;
} VkrBlurParams;

/* GTAO_main / GTAO_filter: UBO GTAOParams (reference src/gtao.hpp:12-18) + push
 * {angle_offset, weight_ratio, use_mis, two_directions, reflections_only} (src/gtao.cpp:101-113);
 * the filter uses znear/zfar only (:211-216). */
typedef struct VkrGtaoParams {
  float normal_mat[16];
  float fovy, aspect, znear, zfar;
  float angle_offset, weight_ratio;
  uint32_t use_mis, two_directions, reflections_only;
} VkrGtaoParams;

/* GTAO_accumulate: UBO AccumConstants + push {clear_history} (reference src/gtao.cpp:300-313). */
typedef struct VkrAccumParams {
  float inverse_camera[16];
  float prev_inverse_camera[16];
  float mvp[16];
  float fovy_aspect_znear_zfar[4];
  uint32_t clear_history;
} VkrAccumParams;

/* DeferedShading: UBO ShaderConstants (reference src/defered_shading.cpp:4-12) + push
 * {min_max_roughness, show_ao} (:68-72).  shadow_mvp is carried but, like the reference
 * shader, never used. */
typedef struct VkrShadingParams {
  float inverse_camera[16];
  float camera[16];
  float shadow_mvp[16];
  float fovy, aspect, znear, zfar;
  float min_max_roughness[2];
  uint32_t show_ao;
} VkrShadingParams;

/* TAA: UBO TAAParams (reference src/taa.cpp:24-30). */
typedef struct VkrTaaParams {
  float inverse_camera[16];
  float prev_inverse_camera[16];
  float fovy_aspect_znear_zfar[4];
} VkrTaaParams;

/* SSAO: UBO SSAOParams as the SHADER declares it (reference shaders/ssao/shader.frag:11-19; std140, so every vec3 of
 * samples[] occupies 16 bytes — samples[i][3] is padding).  The reference's C++ mirror of the block packs glm::vec3 at
 * 12 bytes (src/ssao.cpp:15-22), so its shader reads the sample floats at a 16-byte stride out of a 12-byte-strided
 * array; a caller that wants those very bytes copies its 48 packed floats into samples[0..11][0..3]. */
typedef struct VkrSsaoParams {
  float projection[16];
  float fovy, aspect, znear, zfar;
  float samples[16][4];
} VkrSsaoParams;

/* Legacy SSR: UBO SSRParams (reference src/ssr.hpp:11-17). */
typedef struct VkrSsrParams {
  float normal_mat[16];
  float fovy, aspect, znear, zfar;
} VkrSsrParams;

/* ScreenSpaceTrace: GpuParams of add_main_pass (reference src/screen_trace.cpp:31-39).  random_offset and the random part
 * of angle_offset come from the pass's std::default_random_engine in the reference (:49-54); here they are inputs. */
typedef struct VkrScreenTraceParams {
  float normal_mat[16];
  float random_offset, angle_offset;
  float fovy, aspect, znear, zfar;
} VkrScreenTraceParams;

/* ---- extended launches: array layers and row windows ------------------------------------
 * Every pass entry has an `_ex` twin taking a VkrLaunchExt (NULL = the plain call):
 *  - layers > 1: each VkrImage argument describes layer 0 of an array image whose layers lie
 *    layer_stride[i] bytes apart (i = position of the image among the entry's VkrImage arguments,
 *    0-based; stride 0 = one image shared by every layer, e.g. a LUT), `p` points to `layers`
 *    params structs in HOST memory and params_scratch to device memory of at least
 *    layers * VKR_PARAMS_SCRATCH_STRIDE bytes that stays untouched until the stream passes the call.
 *    Layers ride on grid.z, so hundreds of small views fill the GPU in one launch.  This is the
 *    array_layers / base_layer / layers_count notion of gpu::ImageInfo / gpu::ImageViewRange
 *    (reference src/gpu/resource_info.hpp:11-45) that probe_renderer renders its cube faces into
 *    (src/probe_renderer.cpp:205-235).
 *  - row_begin/row_end: compute only output rows [row_begin, row_end) of the pass's output image
 *    (0,0 = all).  Both must be multiples of VKR_ROW_ALIGN unless row_end is the image height.
 *    Used to split one very large frame into row tiles across GPUs; the caller supplies whatever
 *    rows of the inputs the pass reads beyond its window (halo exchange).  No reference counterpart
 *    (BASELINE.json configs[4]).  vkr_hiz_build_ex rejects row windows (the pyramid is built whole).
 *  - flags (layers > 1 only): VKR_EXT_PREPARE_ONLY converts and uploads the params blocks into params_scratch and
 *    returns without launching; VKR_EXT_PARAMS_PREPARED launches with the blocks an earlier call of the SAME entry
 *    left in params_scratch (no host conversion, no copy: the call is then pure kernel launches and can be captured
 *    into a CUDA graph); first_param: see the struct. */
#define VKR_ROW_ALIGN 16
#define VKR_PARAMS_SCRATCH_STRIDE 256
#define VKR_MAX_IMAGE_ARGS 10
#define VKR_EXT_PREPARE_ONLY 1u
#define VKR_EXT_PARAMS_PREPARED 2u
typedef struct VkrLaunchExt {
  uint32_t layers;
  uint32_t row_begin, row_end;
  uint32_t flags;
  uint64_t layer_stride[VKR_MAX_IMAGE_ARGS];
  void* params_scratch;
  uint32_t first_param;   /* with VKR_EXT_PARAMS_PREPARED: layer z of this launch uses block first_param + z of the prepared set, so a
                           * batch prepared once can be launched as several layer ranges (each range's VkrImage arguments then
                           * describe its first layer).  Independent ranges on separate streams fill each other's tails. */
  uint32_t reserved;
  /* Fused gather (vkr_sssr_blur_ex and vkr_gtao_accumulate_ex only; every other entry refuses a non-zero peer_count): every texel
   * the pass stores into its OUTPUT image is also stored at the same address + peer_delta[k], k < peer_count — the same array in
   * peer-mapped memory of other GPUs (cudaIpc / NVLink), so that the result is "gathered" by the kernel that produces it: no
   * collective, no copy, no SMs taken from the passes (include/vkr_peer.h vkr_peer_epoch_begin / _end order it).  No reference
   * counterpart (the reference is single-GPU); BASELINE.json configs[3]. */
  uint32_t peer_count;
  uint32_t reserved2;
  int64_t peer_delta[7];
} VkrLaunchExt;
#define VKR_MAX_PEERS 7

/* ---- utilities ---------------------------------------------------------------- */
const char* vkr_version(void);
/* Thread-local text of the last failure on this thread ("" if none). */
const char* vkr_last_error(void);
/* Bytes per texel of a VkrFormat (0 if unknown). */
uint32_t vkr_format_texel_size(uint32_t format);
/* Fills width/height/format/mip_count/mip_offset/row_pitch of *img (base untouched) with the
 * layout every pass expects (pitch = align16(width*bpp), mip offsets 256-byte aligned) and
 * returns the allocation size in bytes (0 on bad arguments).  Replaces VMA image creation
 * behind RenderGraph::create_image (reference src/rendergraph/rendergraph.cpp:341-369). */
uint64_t vkr_image_layout(uint32_t width, uint32_t height, uint32_t format, uint32_t mip_count, VkrImage* img);
/* The rule every pass entry applies to its VkrImage arguments: base, mip_offset[i] and row_pitch[i] 16-byte aligned
 * (rows are read as 128-bit vectors), row_pitch[i] >= width_i * texel size, known format, 1 <= mip_count <= VKR_MAX_MIPS.
 * VKR_OK or VKR_ERR_INVALID_ARGUMENT (vkr_last_error() names the field).  No reference counterpart: there a VkImage
 * cannot be mis-described (gpu::ImageInfo, reference src/gpu/resource_info.hpp:11-45). */
int vkr_image_validate(const VkrImage* img);
/* floor(log2(max(w,h))) + 1 — the mip count Gbuffer gives its depth images
 * (reference src/scene_renderer.cpp:13). */
uint32_t vkr_full_mip_count(uint32_t width, uint32_t height);
/* reference src/advanced_ssr.cpp:8-34: 128 x vec4, .x = halton(i+1,2), .y = halton(i+1,3). Host memory. */
void vkr_halton23_seq(uint32_t count, float* out_vec4);

/* Test hook: exhaustive device check of the kernels' 3-instruction unorm decode against the IEEE
 * division; returns the number of mismatching codes (0 expected), negative on a CUDA error. */
long long vkr_selftest_decode(void);

/* ---- Hi-Z: DownsamplePass (reference src/downsample_pass.cpp) ------------------- */
/* Task "DownsampleGbuffer" (:63-90, shaders/advanced_ssr/downsample_gbuffer.frag):
 * depth mip0 -> depth mip1 (2x2 min), normal/velocity of the min-depth texel -> out_*. */
int vkr_downsample_gbuffer(const VkrImage* depth, const VkrImage* normal, const VkrImage* velocity,
                           const VkrImage* out_normal, const VkrImage* out_velocity, void* stream);
/* Tasks "DownsampleDepth" x (L - src_mip - 1) (:107-130, shaders/advanced_ssr/depth_mips.frag):
 * builds mips src_mip+1 .. L-1 from mip src_mip. */
int vkr_depth_mips(const VkrImage* depth, uint32_t src_mip, void* stream);
/* Fused DownsamplePass::run (:133-143): both of the above in one launch + one tail launch. */
int vkr_hiz_build(const VkrImage* depth, const VkrImage* normal, const VkrImage* velocity,
                  const VkrImage* out_normal, const VkrImage* out_velocity, void* stream);

/* ---- AdvancedSSR (reference src/advanced_ssr.cpp) -------------------------------- */
/* Task "SSR_preintegrate" (:95-114, preintegrate.comp) -> R32_SFLOAT LUT. */
int vkr_pdf_preintegrate(const VkrImage* out_pdf, void* stream);
/* Task "BRDF_preintegrate" (:116-136, preintegrate_ssr.comp) -> R16G16_SFLOAT LUT.
 * halton_vec4x128: DEVICE pointer to 128 x float4 (the Halton UBO). */
int vkr_brdf_preintegrate(const float* halton_vec4x128, const VkrImage* out_brdf, void* stream);
/* Task "SSSR_trace" (:147-214, trace.comp). depth = whole D24 image (the callee applies the
 * mips 1..L-1 view); out_occlusion = gtao.raw. */
int vkr_sssr_trace(const VkrTraceParams* p, const VkrImage* depth, const VkrImage* dnormal, const VkrImage* material,
                   const float* halton_vec4x128, const VkrImage* pdf_lut, const VkrImage* out_rays,
                   const VkrImage* out_occlusion, void* stream);
/* Task "SSSR_filter" (:308-369, filter.comp). */
int vkr_sssr_filter(const VkrTraceParams* p, const VkrImage* rays, const VkrImage* depth, const VkrImage* albedo,
                    const VkrImage* normal, const VkrImage* material, const VkrImage* out_reflections, void* stream);
/* Task "SSSR_blur" (:371-438, blur.comp). */
int vkr_sssr_blur(const VkrBlurParams* p, const VkrImage* depth, const VkrImage* normal, const VkrImage* reflections,
                  const VkrImage* material, const VkrImage* history, const VkrImage* dvelocity,
                  const VkrImage* prev_depth, const VkrImage* out_blurred, void* stream);

/* ---- SSSR tile path (reference src/advanced_ssr.cpp:216-302,:440-538; left out of AdvancedSSR::run at :547-550) -------------
 * Buffers are plain device pointers the callee borrows: *_indirect = one VkDispatchIndirectCommand {x, y, z} (12 bytes) whose .x
 * counts the tiles of its list; *_tiles = int32 tile indices (tile = 8x8 block of the half-res `rays` image, index =
 * tile_y * ceil(w/8) + tile_x), capacity >= the number of tiles.  The list ORDER comes out of atomicAdd, as in the reference. */
typedef struct VkrClassifyParams {   /* push constants of SSSR_Classification (src/advanced_ssr.cpp:464-471) */
  int32_t width, height;             /* extent of `rays` (half-res) */
  float max_roughness;
  float glossy_value;
} VkrClassifyParams;
typedef struct VkrRegressionParams { /* push constants of SSSR_Tile_Regression (src/advanced_ssr.cpp:508-518) */
  float camera_to_world[16];         /* the reference passes transpose(normal_mat) */
  float fovy, aspect, znear, zfar;
} VkrRegressionParams;
/* Task "SSSR_Clear" (:440-452): both commands <- {0, 1, 1}. */
int vkr_sssr_clear_indirect(void* reflective_indirect, void* glossy_indirect, void* stream);
/* Task "SSSR_Classification" (:454-495, classification.comp): mean roughness of each tile < glossy_value -> reflective list,
 * else glossy list.  material = the full-res image, sampled at pixel / extent. */
int vkr_sssr_classification(const VkrClassifyParams* p, const VkrImage* material, int32_t* reflective_tiles, int32_t* glossy_tiles,
                            void* reflective_indirect, void* glossy_indirect, void* stream);
/* One of the two vkCmdDispatchIndirect of task "SSSR_trace" (:216-302, trace_indirect.comp): reflection_type 0 = mirror list
 * (50 cells from view level 0 + thickness test), 1 = glossy list (25 cells from view level 1).  The tile count is read on the
 * device from indirect->x (clamped to max_tiles, the list capacity): no host synchronisation.  p->render_flags is not read. */
int vkr_sssr_trace_indirect(const VkrTraceParams* p, uint32_t reflection_type, const VkrImage* depth, const VkrImage* dnormal,
                            const VkrImage* material, const float* halton_vec4x128, const int32_t* tiles, const void* indirect,
                            uint32_t max_tiles, const VkrImage* out_rays, void* stream);
/* Task "SSSR_Tile_Regression" (:497-538, regression.comp): per half-res 8x8 tile the plane n with n . x = 1 through the tile's
 * camera-relative world positions (.xyz) and the mean squared residual (.w) -> out_planes R32G32B32A32_SFLOAT. */
int vkr_sssr_tile_regression(const VkrRegressionParams* p, const VkrImage* depth, const VkrImage* out_planes, void* stream);
/* The tile path of the trace in ONE call that allocates nothing: SSSR_Clear -> SSSR_Classification -> both indirect SSSR_trace
 * dispatches, with the two tile lists, the two indirect commands and the sRGB decode table in a file-scope device array inside
 * the library instead of caller-allocated buffers.  NO COUNTERPART IN THE REFERENCE: added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).
 *  - Calls no cudaMalloc / cudaFree / allocating library routine.  The FIRST call on a device does the set-up (symbol address
 *    lookup + a synchronous 1 KB table upload): make that call outside stream capture.
 *  - Limit: ceil(w/8) * ceil(h/8) <= VKR_SSSR_STATIC_MAX_TILES for the w x h `out_rays` image (65 536 tiles: up to 2048 x 2048,
 *    or the half-res rays of a 3840 x 2160 frame); larger extents return VKR_ERR_UNSUPPORTED with a message naming the limit.
 *    Every GPU that loads the library holds the array (VKR_SSSR_STATIC_SCRATCH_BYTES, about 513 KB) until exit.
 *  - Every call shares that array; the counters are cleared at the start of every call.  Calls that OVERLAP on the GPU (different
 *    streams without an event between them, or several host threads on one device) race on it: serialise them.
 *  - Inputs and output as vkr_sssr_classification + vkr_sssr_trace_indirect: classification runs over the extent of `out_rays`
 *    with p->max_roughness and `glossy_value`; every pixel of `out_rays` is written (each tile is in exactly one list); the
 *    result equals, bit for bit, the separate entries' (tests/test_zz_no_alloc.py).  halton_vec4x128: DEVICE memory. */
#define VKR_SSSR_STATIC_MAX_TILES 65536u
#define VKR_SSSR_STATIC_SCRATCH_BYTES ((64u + 2u * VKR_SSSR_STATIC_MAX_TILES + 256u) * 4u)
int vkr_sssr_trace_tiles_static(const VkrTraceParams* p, float glossy_value, const VkrImage* depth, const VkrImage* dnormal,
                                const VkrImage* material, const float* halton_vec4x128, const VkrImage* out_rays, void* stream);

/* ---- capture path (reference src/util_passes.cpp:152-180, src/image_readback.cpp:67-110, src/main.cpp:152-176) ------------ */
/* Task "CopyImage" (blit_image): vkCmdBlitImage, VK_FILTER_LINEAR, all of src mip 0 onto all of dst mip 0 — destination texel p
 * takes the bilinear, clamp-to-edge sample of src at (p + 0.5) / dst_extent, converted to dst's format.  src: R16G16B16A16_SFLOAT,
 * R8G8B8A8_SRGB or R8G8B8A8_UNORM; dst: R8G8B8A8_SRGB or R8G8B8A8_UNORM (the reference blits the TAA target into its sRGB
 * read-back image, src/main.cpp:220-223,:394). */
int vkr_blit_image(const VkrImage* src, const VkrImage* dst, void* stream);
/* Task "ImageRead" (ReadBackSystem::read_image): one mip level into a tightly packed DEVICE buffer of width * height * texel
 * bytes (bufferRowLength = 0), ready for one contiguous device-to-host copy.  mode applies what the reference's callbacks do
 * to the bytes on the host: DEPTH24 = 32-bit words & 0xFFFFFF (get_depth_cb / get_depth_png_cb), OPAQUE_ALPHA = byte 3 of every
 * texel <- 255 (get_rgba_cb); both need 4-byte texels. */
typedef enum VkrReadbackMode { VKR_READBACK_RAW = 0, VKR_READBACK_DEPTH24 = 1, VKR_READBACK_OPAQUE_ALPHA = 2 } VkrReadbackMode;
int vkr_readback_pack(const VkrImage* image, uint32_t mip, uint32_t mode, void* out_tight, void* stream);

/* ---- GTAO (reference src/gtao.cpp) ------------------------------------------------ */
/* Task "GTAO_main" (:84-148, gtao/main.comp); raw is read-modify-write. depth_lod = GTAO::depth_lod (:23). */
int vkr_gtao_main(const VkrGtaoParams* p, const VkrImage* depth, const VkrImage* normal, const VkrImage* material,
                  const VkrImage* pdf_lut, const VkrImage* raw, uint32_t depth_lod, void* stream);
/* Task "GTAO_filter" (:198-239, gtao/filter.comp). */
int vkr_gtao_filter(const VkrGtaoParams* p, const VkrImage* depth, const VkrImage* raw, const VkrImage* filtered,
                    uint32_t depth_lod, void* stream);
/* Task "GTAO_accumulate" (:286-347, gtao/accum.comp). */
int vkr_gtao_accumulate(const VkrAccumParams* p, const VkrImage* depth, const VkrImage* prev_depth,
                        const VkrImage* filtered, const VkrImage* accumulated, const VkrImage* dvelocity,
                        const VkrImage* history, uint32_t depth_lod, void* stream);

/* ---- DeferedShadingPass / TAA ------------------------------------------------------ */
/* Task "DeferedShading" (reference src/defered_shading.cpp:47-118, defered_shading/shader.frag). */
int vkr_deferred_shading(const VkrShadingParams* p, const VkrImage* albedo, const VkrImage* normal,
                         const VkrImage* material, const VkrImage* depth, const VkrImage* occlusion,
                         const VkrImage* brdf_lut, const VkrImage* reflections, const VkrImage* out_color, void* stream);
/* Task "TAA" (reference src/taa.cpp:14-63, taa/resolve.comp). */
int vkr_taa_resolve(const VkrTaaParams* p, const VkrImage* history, const VkrImage* prev_depth, const VkrImage* depth,
                    const VkrImage* velocity, const VkrImage* color, const VkrImage* out_target, void* stream);

/* ---- earlier passes kept beside the live ones (north_star: "GTAO/SSAO", "SSR/advanced_ssr") ------ */
/* Task "SSAO" (reference src/ssao.cpp:54-96 SSAOPass::draw, ssao/shader.frag): depth mip 0, out R8_UNORM (:8-11).
 * With layers > 1 every layer uses p[0].samples (one sample set per pass object, :34-52). */
int vkr_ssao(const VkrSsaoParams* p, const VkrImage* depth, const VkrImage* out_occlusion, void* stream);
/* Task "SSR" (reference src/ssr.cpp:10-73 add_ssr_pass, ssr/shader.frag + include/screen_trace.glsl:51-100).
 * depth = whole D24 image with its min-pyramid; frame = the lit colour image (R8G8B8A8_SRGB or R16G16B16A16_SFLOAT);
 * out R8G8B8A8_UNORM (:5-8). */
int vkr_ssr_legacy(const VkrSsrParams* p, const VkrImage* depth, const VkrImage* normal, const VkrImage* frame,
                   const VkrImage* material, const VkrImage* out_reflection, void* stream);

/* Tasks "ScreenTrace", "ScreenTraceFilter", "ScreenTraceAccumulate" (reference src/screen_trace.cpp:23-99, :101-139,
 * :141-180; shaders/screen_trace/{trace,filter,accumulate}.comp).  depth / prev_depth: mip 0 is read; color = the lit frame
 * (R8G8B8A8_SRGB or R16G16B16A16_SFLOAT); raw / filtered / accumulated are R16G16B16A16_SFLOAT (:4-9), accumulated is
 * read-modify-write.  Like the reference's dispatches, only floor(W/8) x floor(H/8) tiles (main) and
 * floor(W/8)*8 x floor(H/4)*4 pixels (filter, accumulate) are written.  The tile exchange of trace.comp:305-306 (no
 * barrier() in the reference) is executed as if barriered; sky pixels publish "no hit" (csrc/screen_trace.cu header). */
int vkr_screen_trace_main(const VkrScreenTraceParams* p, const VkrImage* depth, const VkrImage* normal, const VkrImage* color,
                          const VkrImage* material, const VkrImage* out_raw, void* stream);
int vkr_screen_trace_filter(float znear, float zfar, const VkrImage* raw, const VkrImage* depth, const VkrImage* out_filtered, void* stream);
int vkr_screen_trace_accumulate(const float* fovy_aspect_znear_zfar, const VkrImage* depth, const VkrImage* prev_depth,
                                const VkrImage* filtered, const VkrImage* accumulated, void* stream);

/* GTAO variants beside the live compute path.  p->angle_offset is the push constant; the MIS fields of VkrGtaoParams are not
 * read.  `ext` (may be NULL) selects a row window; array layers are not supported by these four.
 * Task "GTAO" (reference src/gtao.cpp:349-413 add_main_pass_graphics, gtao/main.frag): every pixel of raw, 1 slice x 20 taps;
 * writes (occlusion, 0, 0, 1) — the reference's float fragment output leaves .gba of the RGBA16F attachment undefined. */
int vkr_gtao_main_graphics(const VkrGtaoParams* p, const VkrImage* depth, const VkrImage* normal, const VkrImage* raw, uint32_t depth_lod,
                           const VkrLaunchExt* ext, void* stream);
/* Task "GTAO_deinterleave" (:445-472, gtao_opt/deinterleave.comp): out_array = R32_SFLOAT, 4^pattern_n layers layer_stride bytes
 * apart.  As in the reference the dispatch covers the OUTPUT extent, so only that corner of the depth level is scattered. */
int vkr_gtao_deinterleave(const VkrImage* depth, uint32_t depth_lod, const VkrImage* out_array, uint64_t layer_stride, int pattern_n, void* stream);
/* Task "GTAO_deinterleaved" (:474-526, gtao_opt/main_deinterleaved.comp) for ONE array layer; the reference loops over the
 * layers of its output image (one).  Invocation i of the dispatch over raw's extent writes pixel 2^n * i + layer offset. */
int vkr_gtao_main_deinterleaved(const VkrGtaoParams* p, const VkrImage* depth_array, uint64_t layer_stride, const VkrImage* normal,
                                const VkrImage* raw, int pattern_n, uint32_t layer, const VkrLaunchExt* ext, void* stream);
/* Task "GTAO_reproject" (:241-284, gtao/reproject.comp in its STATIC_REPROJECT mode: the matrix of GTAOReprojection is unused).
 * gtao = filtered, prev_gtao = prev_frame, out = output, all R16_SFLOAT. */
int vkr_gtao_reproject(const float* fovy_aspect_znear_zfar, const VkrImage* depth, const VkrImage* prev_depth, uint32_t depth_lod,
                       const VkrImage* gtao, const VkrImage* prev_gtao, const VkrImage* out, const VkrLaunchExt* ext, void* stream);

/* ---- _ex twins (see VkrLaunchExt) ---------------------------------------------------- */
int vkr_hiz_build_ex(const VkrImage* depth, const VkrImage* normal, const VkrImage* velocity, const VkrImage* out_normal,
                     const VkrImage* out_velocity, const VkrLaunchExt* ext, void* stream);
int vkr_sssr_trace_ex(const VkrTraceParams* p, const VkrImage* depth, const VkrImage* dnormal, const VkrImage* material,
                      const float* halton_vec4x128, const VkrImage* pdf_lut, const VkrImage* out_rays,
                      const VkrImage* out_occlusion, const VkrLaunchExt* ext, void* stream);
int vkr_sssr_filter_ex(const VkrTraceParams* p, const VkrImage* rays, const VkrImage* depth, const VkrImage* albedo,
                       const VkrImage* normal, const VkrImage* material, const VkrImage* out_reflections,
                       const VkrLaunchExt* ext, void* stream);
int vkr_sssr_blur_ex(const VkrBlurParams* p, const VkrImage* depth, const VkrImage* normal, const VkrImage* reflections,
                     const VkrImage* material, const VkrImage* history, const VkrImage* dvelocity,
                     const VkrImage* prev_depth, const VkrImage* out_blurred, const VkrLaunchExt* ext, void* stream);
int vkr_gtao_main_ex(const VkrGtaoParams* p, const VkrImage* depth, const VkrImage* normal, const VkrImage* material,
                     const VkrImage* pdf_lut, const VkrImage* raw, uint32_t depth_lod, const VkrLaunchExt* ext, void* stream);
int vkr_gtao_filter_ex(const VkrGtaoParams* p, const VkrImage* depth, const VkrImage* raw, const VkrImage* filtered,
                       uint32_t depth_lod, const VkrLaunchExt* ext, void* stream);
int vkr_gtao_accumulate_ex(const VkrAccumParams* p, const VkrImage* depth, const VkrImage* prev_depth,
                           const VkrImage* filtered, const VkrImage* accumulated, const VkrImage* dvelocity,
                           const VkrImage* history, uint32_t depth_lod, const VkrLaunchExt* ext, void* stream);
int vkr_deferred_shading_ex(const VkrShadingParams* p, const VkrImage* albedo, const VkrImage* normal,
                            const VkrImage* material, const VkrImage* depth, const VkrImage* occlusion,
                            const VkrImage* brdf_lut, const VkrImage* reflections, const VkrImage* out_color,
                            const VkrLaunchExt* ext, void* stream);
int vkr_taa_resolve_ex(const VkrTaaParams* p, const VkrImage* history, const VkrImage* prev_depth, const VkrImage* depth,
                       const VkrImage* velocity, const VkrImage* color, const VkrImage* out_target,
                       const VkrLaunchExt* ext, void* stream);
int vkr_ssao_ex(const VkrSsaoParams* p, const VkrImage* depth, const VkrImage* out_occlusion, const VkrLaunchExt* ext, void* stream);
int vkr_ssr_legacy_ex(const VkrSsrParams* p, const VkrImage* depth, const VkrImage* normal, const VkrImage* frame,
                      const VkrImage* material, const VkrImage* out_reflection, const VkrLaunchExt* ext, void* stream);
int vkr_screen_trace_main_ex(const VkrScreenTraceParams* p, const VkrImage* depth, const VkrImage* normal, const VkrImage* color,
                             const VkrImage* material, const VkrImage* out_raw, const VkrLaunchExt* ext, void* stream);
int vkr_screen_trace_filter_ex(float znear, float zfar, const VkrImage* raw, const VkrImage* depth, const VkrImage* out_filtered,
                               const VkrLaunchExt* ext, void* stream);
int vkr_screen_trace_accumulate_ex(const float* fovy_aspect_znear_zfar, const VkrImage* depth, const VkrImage* prev_depth,
                                   const VkrImage* filtered, const VkrImage* accumulated, const VkrLaunchExt* ext, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* VKR_PASSES_H_INCLUDED */
