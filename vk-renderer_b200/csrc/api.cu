// api.cu — C-ABI utilities: errors, image layout, host-built constant tables.
#include <cmath>
#include <cstdarg>
#include <cstdio>
#include <cstring>
#include <mutex>
#include "common.cuh"

namespace vkr {

static thread_local char g_err[512] = "";
static thread_local char g_image_problem[192] = "";

void note_image_problem(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_image_problem, sizeof(g_image_problem), fmt, ap);
  va_end(ap);
}

void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  int n = vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
  if (g_image_problem[0] && n >= 0 && n < (int)sizeof(g_err) - 4) {      // image_ok() said why: keep it with the caller's message
    snprintf(g_err + n, sizeof(g_err) - n, " [image layout: %s]", g_image_problem);
    g_image_problem[0] = 0;
  }
}

// VKR_HOST_ONLY (oracle/Makefile, target libvkr_hostonly.so, compiled by g++): the layout / table / producer-twin helpers
// without any device code or CUDA runtime call, so that bench.py's CPU arm can build its inputs without loading the product
// library.  Never defined in the product build.
CamK make_cam(float fovy, float aspect, float znear, float zfar) {
  CamK c;
  c.aspect = aspect; c.znear = znear; c.zfar = zfar;
  c.tg = std::tan(fovy / 2);  // gbuffer_encode.glsl:60,:76 — hoisted, same float the shader would form
  return c;
}

#ifndef VKR_HOST_ONLY
int check_launch(const char* what) {
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) {
    set_error("%s: %s", what, cudaGetErrorString(e));
    return VKR_ERR_CUDA;
  }
  return VKR_OK;
}

int resolve_ext(const VkrLaunchExt* e, uint32_t out_h, int tile_h, ExtInfo* o, const char* what, bool peers_ok) {
  o->e = e;
  if (e && e->peer_count && (!peers_ok || e->peer_count > VKR_MAX_PEERS)) {
    set_error("%s: peer_count = %u (fused gather) is %s", what, e->peer_count, peers_ok ? "beyond VKR_MAX_PEERS" : "only supported by sssr_blur_ex and gtao_accumulate_ex");
    return VKR_ERR_UNSUPPORTED;
  }
  o->layers = (e && e->layers > 1) ? (int)e->layers : 1;
  const int total = (int)((out_h + tile_h - 1) / tile_h);
  o->yb0 = 0; o->yblocks = total;
  if (e && (e->row_begin != 0 || e->row_end != 0)) {
    const uint32_t y0 = e->row_begin, y1 = e->row_end < out_h ? e->row_end : out_h;
    if (y0 >= y1 || (y0 % VKR_ROW_ALIGN) != 0 || (y1 != out_h && (y1 % VKR_ROW_ALIGN) != 0) || (VKR_ROW_ALIGN % tile_h) != 0) {
      set_error("%s: row window [%u, %u) must be non-empty and aligned to %d rows (or end at the image height %u)", what, e->row_begin, e->row_end, VKR_ROW_ALIGN, out_h);
      return VKR_ERR_INVALID_ARGUMENT;
    }
    o->yb0 = (int)(y0 / tile_h);
    o->yblocks = (int)((y1 - y0 + tile_h - 1) / tile_h);
  }
  if (o->layers > 1 && !e->params_scratch) {
    set_error("%s: layers > 1 needs params_scratch (device memory, layers * VKR_PARAMS_SCRATCH_STRIDE bytes)", what);
    return VKR_ERR_INVALID_ARGUMENT;
  }
  if (e && e->first_param && !(o->layers > 1 && (e->flags & VKR_EXT_PARAMS_PREPARED))) {
    set_error("%s: first_param needs layers > 1 and VKR_EXT_PARAMS_PREPARED (it selects blocks of a set prepared earlier)", what);
    return VKR_ERR_INVALID_ARGUMENT;
  }
  if (o->layers > 65535) { set_error("%s: at most 65535 layers per launch", what); return VKR_ERR_INVALID_ARGUMENT; }
  return VKR_OK;
}

int upload_params(const void* host, size_t elem, int count, void* scratch, cudaStream_t st, const char* what) {
  if (elem > VKR_PARAMS_SCRATCH_STRIDE) { set_error("%s: params block larger than the scratch stride", what); return VKR_ERR_UNSUPPORTED; }
  // pageable source: the runtime stages it before returning, so `host` may be a temporary
  cudaError_t e = cudaMemcpyAsync(scratch, host, elem * (size_t)count, cudaMemcpyHostToDevice, st);
  if (e != cudaSuccess) { set_error("%s: params upload: %s", what, cudaGetErrorString(e)); return VKR_ERR_CUDA; }
  return VKR_OK;
}

// sRGB EOTF table, built in double precision on the host, one copy per device.
static std::mutex g_tab_mutex;
static float* g_srgb_dev[64] = {nullptr};

const float* srgb_decode_table_device() {
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return nullptr;
  std::lock_guard<std::mutex> lock(g_tab_mutex);
  if (!g_srgb_dev[dev]) {
    float host[256];
    for (int i = 0; i < 256; i++) {
      double c = i / 255.0;
      host[i] = (float)((c <= 0.04045) ? c / 12.92 : std::pow((c + 0.055) / 1.055, 2.4));
    }
    float* d = nullptr;
    if (cudaMalloc(&d, sizeof(host)) != cudaSuccess) return nullptr;
    if (cudaMemcpy(d, host, sizeof(host), cudaMemcpyHostToDevice) != cudaSuccess) { cudaFree(d); return nullptr; }
    g_srgb_dev[dev] = d;
  }
  return g_srgb_dev[dev];
}

// exhaustive device check of the 3-instruction unorm decodes against the IEEE division
__global__ void selftest_decode_kernel(unsigned long long* mismatches) {
  unsigned k = blockIdx.x * blockDim.x + threadIdx.x;   // 2^24 threads
  unsigned bad = 0;
  bad += d24_to_float(k) != __fdiv_rn((float)k, 16777215.0f);
  if (k < 65536u) bad += unorm16_to_float(k) != __fdiv_rn((float)k, 65535.0f);
  if (k < 256u) bad += unorm8_to_float(k) != __fdiv_rn((float)k, 255.0f);
  if (bad) atomicAdd(mismatches, (unsigned long long)bad);
}

#endif  // !VKR_HOST_ONLY

}  // namespace vkr

extern "C" {

#ifndef VKR_HOST_ONLY
/* Test hook: number of inputs for which the fast unorm decode differs from __fdiv_rn (must be 0);
 * negative on a CUDA error. */
long long vkr_selftest_decode(void) {
  unsigned long long* d = nullptr;
  unsigned long long h = 0;
  if (cudaMalloc(&d, sizeof(h)) != cudaSuccess) return -1;
  cudaMemset(d, 0, sizeof(h));
  vkr::selftest_decode_kernel<<<(1u << 24) / 256, 256>>>(d);
  cudaError_t e = cudaMemcpy(&h, d, sizeof(h), cudaMemcpyDeviceToHost);
  cudaFree(d);
  if (e != cudaSuccess) { vkr::set_error("selftest_decode: %s", cudaGetErrorString(e)); return -1; }
  return (long long)h;
}
#endif  // !VKR_HOST_ONLY

const char* vkr_version(void) { return "vkr_b200 0.1 (sm_100a)"; }
const char* vkr_last_error(void) { return vkr::g_err; }

uint32_t vkr_format_texel_size(uint32_t f) {
  switch (f) {
    case VKR_FORMAT_R8_UNORM: return 1;
    case VKR_FORMAT_R16_UNORM: case VKR_FORMAT_R16_SFLOAT: return 2;
    case VKR_FORMAT_R8G8B8A8_UNORM: case VKR_FORMAT_R8G8B8A8_SRGB: case VKR_FORMAT_R16G16_UNORM:
    case VKR_FORMAT_R16G16_SFLOAT: case VKR_FORMAT_R32_SFLOAT: case VKR_FORMAT_D24_UNORM_S8_UINT: return 4;
    case VKR_FORMAT_R16G16B16A16_UNORM: case VKR_FORMAT_R16G16B16A16_SFLOAT: return 8;
    case VKR_FORMAT_R32G32B32A32_SFLOAT: return 16;
  }
  return 0;
}

int vkr_image_validate(const VkrImage* img) {
  if (!img || !img->base || !img->width || !img->height || img->mip_count < 1 || img->mip_count > VKR_MAX_MIPS) {
    vkr::set_error("image_validate: null image, null base, zero extent or mip_count outside [1, %d]", VKR_MAX_MIPS);
    return VKR_ERR_INVALID_ARGUMENT;
  }
  if (!vkr::image_layout_ok(img)) { vkr::set_error("image_validate: layout refused"); return VKR_ERR_INVALID_ARGUMENT; }
  return VKR_OK;
}

uint32_t vkr_full_mip_count(uint32_t w, uint32_t h) {
  uint32_t m = w > h ? w : h, n = 0;
  while (m) { n++; m >>= 1; }
  return n;
}

uint64_t vkr_image_layout(uint32_t width, uint32_t height, uint32_t format, uint32_t mip_count, VkrImage* img) {
  uint32_t bpp = vkr_format_texel_size(format);
  if (!img || !bpp || !width || !height || mip_count < 1 || mip_count > VKR_MAX_MIPS) return 0;
  if (mip_count > vkr_full_mip_count(width, height)) return 0;
  img->width = width; img->height = height; img->format = format; img->mip_count = mip_count;
  uint64_t off = 0;
  for (uint32_t i = 0; i < VKR_MAX_MIPS; i++) { img->mip_offset[i] = 0; img->row_pitch[i] = 0; }
  for (uint32_t i = 0; i < mip_count; i++) {
    uint32_t w = vkr::mip_dim(width, i), h = vkr::mip_dim(height, i);
    uint32_t pitch = (w * bpp + 15u) & ~15u;
    img->mip_offset[i] = off;
    img->row_pitch[i] = pitch;
    off += (uint64_t)pitch * h;
    off = (off + 255u) & ~(uint64_t)255u;
  }
  return off;
}

// reference src/advanced_ssr.cpp:8-34 (radical inverse evaluated in fp32, index + 1)
static float halton_elem(uint32_t index, uint32_t base) {
  float f = 1, r = 0;
  uint32_t current = index;
  do {
    f = f / base;
    r = r + f * (current % base);
    current = (uint32_t)std::floor(current / float(base));
  } while (current > 0);
  return r;
}

void vkr_halton23_seq(uint32_t count, float* out) {
  for (uint32_t i = 0; i < count; i++) {
    out[4 * i] = halton_elem(i + 1, 2);
    out[4 * i + 1] = halton_elem(i + 1, 3);
    out[4 * i + 2] = 0.f; out[4 * i + 3] = 0.f;
  }
}

}  // extern "C"
