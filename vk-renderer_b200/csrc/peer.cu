// peer.cu — IPC-mapped peer memory, halo-row pushes and stream-ordered flags for row-tiled frames (include/vkr_peer.h).
#include "common.cuh"
#include "../../include/vkr_peer.h"

namespace vkr {

__device__ unsigned int g_wait_failed = 0;

__global__ void peer_signal_kernel(uint32_t* flag, uint32_t value) {
  __threadfence_system();                                            // everything this stream wrote before is visible first
  asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(flag), "r"(value) : "memory");
}

__global__ void peer_wait_kernel(const uint32_t* flag, uint32_t value, unsigned long long timeout_ns) {
  unsigned long long t0;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t0));
  for (;;) {
    uint32_t v;
    asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(flag) : "memory");
    if ((int32_t)(v - value) >= 0) return;                           // epochs wrap
    __nanosleep(200);
    unsigned long long t1;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t1));
    if (t1 - t0 > timeout_ns) { atomicAdd(&g_wait_failed, 1u); return; }
  }
}

// Epoch protocol of the fused gather (VkrLaunchExt.peer_delta): both kernels take the epoch from a device word, so a frame that
// contains them can be captured into a CUDA graph once and replayed.
//   ctrl[0]   this rank's epoch (number of frames it has finished writing)
//   flags[r]  the last epoch rank r has finished writing INTO THIS RANK's arrays (stored by r's epoch_end)
// begin: the frame about to run writes epoch e = ctrl[0] + 1 into buffer e & 1 of every rank, whose last tenant was epoch e - 2; a
// peer is done reading that once it has finished ITS epoch e - 1 (the consumer rule: consume epoch k before finishing k + 1).
__global__ void peer_epoch_begin_kernel(const uint32_t* ctrl, const uint32_t* flags, uint32_t count, uint32_t self, unsigned long long timeout_ns) {
  const uint32_t i = threadIdx.x;
  if (i >= count || i == self) return;
  const uint32_t need = ctrl[0];                                      // e - 1
  unsigned long long t0;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t0));
  for (;;) {
    uint32_t v;
    asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(flags + i) : "memory");
    if ((int32_t)(v - need) >= 0) return;
    __nanosleep(200);
    unsigned long long t1;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t1));
    if (t1 - t0 > timeout_ns) { atomicAdd(&g_wait_failed, 1u); return; }
  }
}
struct PeerFlags { uint32_t* p[8]; };
// end: everything this stream stored before (the passes' stores into the peers' arrays included) becomes visible, then the new
// epoch is released into every peer's flag word for this rank.
__global__ void peer_epoch_end_kernel(uint32_t* ctrl, PeerFlags remote, uint32_t count) {
  const uint32_t e = ctrl[0] + 1u;
  __threadfence_system();
  if (threadIdx.x == 0) ctrl[0] = e;
  if (threadIdx.x < count && remote.p[threadIdx.x]) asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(remote.p[threadIdx.x]), "r"(e) : "memory");
}

static int cuda_rc(cudaError_t e, const char* what) {
  if (e == cudaSuccess) return VKR_OK;
  set_error("%s: %s", what, cudaGetErrorString(e));
  return VKR_ERR_CUDA;
}

}  // namespace vkr
using namespace vkr;

extern "C" int vkr_peer_alloc(uint64_t bytes, void** dptr) {
  if (!dptr || !bytes) { set_error("peer_alloc: bad arguments"); return VKR_ERR_INVALID_ARGUMENT; }
  int rc = cuda_rc(cudaMalloc(dptr, bytes), "peer_alloc");
  if (rc) return rc;
  return cuda_rc(cudaMemset(*dptr, 0, bytes), "peer_alloc memset");
}

extern "C" int vkr_peer_free(void* dptr) { return cuda_rc(cudaFree(dptr), "peer_free"); }

extern "C" int vkr_peer_export(void* dptr, VkrIpcHandle* out) {
  static_assert(sizeof(cudaIpcMemHandle_t) <= sizeof(VkrIpcHandle), "handle size");
  if (!dptr || !out) { set_error("peer_export: bad arguments"); return VKR_ERR_INVALID_ARGUMENT; }
  cudaIpcMemHandle_t h;
  int rc = cuda_rc(cudaIpcGetMemHandle(&h, dptr), "peer_export");
  if (rc) return rc;
  memset(out, 0, sizeof(*out));
  memcpy(out->bytes, &h, sizeof(h));
  return VKR_OK;
}

extern "C" int vkr_peer_open(const VkrIpcHandle* handle, void** peer_ptr) {
  if (!handle || !peer_ptr) { set_error("peer_open: bad arguments"); return VKR_ERR_INVALID_ARGUMENT; }
  cudaIpcMemHandle_t h;
  memcpy(&h, handle->bytes, sizeof(h));
  return cuda_rc(cudaIpcOpenMemHandle(peer_ptr, h, cudaIpcMemLazyEnablePeerAccess), "peer_open");
}

extern "C" int vkr_peer_close(void* peer_ptr) { return cuda_rc(cudaIpcCloseMemHandle(peer_ptr), "peer_close"); }

extern "C" int vkr_peer_push_rows(const VkrImage* img, uint32_t lod, uint32_t row_begin, uint32_t row_end, void* peer_image_base, void* stream) {
  if (!img || !img->base || !peer_image_base || lod >= img->mip_count) { set_error("peer_push_rows: bad arguments"); return VKR_ERR_INVALID_ARGUMENT; }
  const uint32_t h = (uint32_t)mip_dim(img->height, lod);
  if (row_end > h) row_end = h;
  if (row_begin >= row_end) return VKR_OK;
  const size_t off = (size_t)img->mip_offset[lod] + (size_t)row_begin * img->row_pitch[lod];
  const size_t bytes = (size_t)(row_end - row_begin) * img->row_pitch[lod];       // full-width rows are contiguous
  return cuda_rc(cudaMemcpyAsync((unsigned char*)peer_image_base + off, (const unsigned char*)img->base + off, bytes, cudaMemcpyDefault, (cudaStream_t)stream),
                 "peer_push_rows");
}

extern "C" int vkr_peer_push_2d(const void* src, uint64_t src_pitch, void* peer_dst, uint64_t dst_pitch, uint64_t width_bytes, uint64_t rows, void* stream) {
  if (!src || !peer_dst || src_pitch < width_bytes || dst_pitch < width_bytes) { set_error("peer_push_2d: bad arguments"); return VKR_ERR_INVALID_ARGUMENT; }
  if (!width_bytes || !rows) return VKR_OK;
  if (src_pitch == width_bytes && dst_pitch == width_bytes)      // one contiguous block: a single DMA descriptor
    return cuda_rc(cudaMemcpyAsync(peer_dst, src, (size_t)(width_bytes * rows), cudaMemcpyDefault, (cudaStream_t)stream), "peer_push_2d");
  return cuda_rc(cudaMemcpy2DAsync(peer_dst, (size_t)dst_pitch, src, (size_t)src_pitch, (size_t)width_bytes, (size_t)rows, cudaMemcpyDefault, (cudaStream_t)stream),
                 "peer_push_2d");
}

extern "C" int vkr_peer_signal(uint32_t* flag, uint32_t value, void* stream) {
  if (!flag) { set_error("peer_signal: null flag"); return VKR_ERR_INVALID_ARGUMENT; }
  peer_signal_kernel<<<1, 1, 0, (cudaStream_t)stream>>>(flag, value);
  return check_launch("peer_signal_kernel");
}

extern "C" int vkr_peer_wait(const uint32_t* flag, uint32_t value, uint32_t timeout_ms, void* stream) {
  if (!flag) { set_error("peer_wait: null flag"); return VKR_ERR_INVALID_ARGUMENT; }
  peer_wait_kernel<<<1, 1, 0, (cudaStream_t)stream>>>(flag, value, (unsigned long long)timeout_ms * 1000000ull);
  return check_launch("peer_wait_kernel");
}

extern "C" int vkr_peer_epoch_begin(const uint32_t* ctrl, const uint32_t* flags, uint32_t count, uint32_t self, uint32_t timeout_ms, void* stream) {
  if (!ctrl || !flags || count == 0 || count > 8 || self >= count) { set_error("peer_epoch_begin: bad arguments (at most 8 ranks)"); return VKR_ERR_INVALID_ARGUMENT; }
  peer_epoch_begin_kernel<<<1, 32, 0, (cudaStream_t)stream>>>(ctrl, flags, count, self, (unsigned long long)timeout_ms * 1000000ull);
  return check_launch("peer_epoch_begin_kernel");
}

extern "C" int vkr_peer_epoch_end(uint32_t* ctrl, uint32_t* const* remote_flags, uint32_t count, void* stream) {
  if (!ctrl || !remote_flags || count == 0 || count > 8) { set_error("peer_epoch_end: bad arguments (at most 8 ranks)"); return VKR_ERR_INVALID_ARGUMENT; }
  PeerFlags f;
  for (uint32_t i = 0; i < 8; i++) f.p[i] = i < count ? remote_flags[i] : nullptr;
  peer_epoch_end_kernel<<<1, 32, 0, (cudaStream_t)stream>>>(ctrl, f, count);
  return check_launch("peer_epoch_end_kernel");
}

extern "C" int vkr_peer_wait_failed(void) {
  unsigned int h = 0, z = 0;
  if (cudaMemcpyFromSymbol(&h, g_wait_failed, sizeof(h)) != cudaSuccess) return -1;
  cudaMemcpyToSymbol(g_wait_failed, &z, sizeof(z));
  return (int)h;
}
