/* vkr_peer.h — peer-memory plumbing for row-tiled frames (BASELINE.json configs[4]: one very large frame split into
 * row tiles across the GPUs of one box, halo rows exchanged over NVLink P2P).
 *
 * No reference counterpart: vk-renderer is single-GPU (src/gpu/driver.cpp:99 picks the first discrete device).  One
 * process per GPU; every rank allocates the SAME slab layout with vkr_peer_alloc, exports it (cudaIpc), opens its
 * neighbours' slabs and then writes halo rows straight into the neighbour's copy of an image (a peer store over
 * NVLink, no staging, no collective).  Ordering between ranks is carried by 32-bit flags inside the slabs:
 *   producer:  vkr_peer_push_rows(...) ; vkr_peer_signal(neighbour_flag, epoch)      (stream-ordered)
 *   consumer:  vkr_peer_wait(my_flag, epoch)  ; <kernel that reads the halo>          (stream-ordered)
 * Nothing here synchronises the host.  All calls return 0 / negative VkrStatus (vkr_last_error() has the text).
 */
#ifndef VKR_PEER_H_INCLUDED
#define VKR_PEER_H_INCLUDED
#include "vkr_passes.h"

#ifdef __cplusplus
extern "C" {
#endif

typedef struct VkrIpcHandle { unsigned char bytes[64]; } VkrIpcHandle;

/* Device allocation that can be exported to other processes (zero-initialised). */
int vkr_peer_alloc(uint64_t bytes, void** dptr);
int vkr_peer_free(void* dptr);
int vkr_peer_export(void* dptr, VkrIpcHandle* out);
/* Maps another process's exported allocation into this process (peer access over NVLink when on another GPU). */
int vkr_peer_open(const VkrIpcHandle* handle, void** peer_ptr);
int vkr_peer_close(void* peer_ptr);

/* Copies rows [row_begin, row_end) of mip `lod` of the local image `img` into the same rows of the peer's copy of that
 * image.  `peer_image_base` is the address of the peer image's allocation base in THIS process (peer slab + the
 * image's offset).  Rows are clamped to the mip's height; an empty range is a no-op.  Asynchronous on `stream`. */
int vkr_peer_push_rows(const VkrImage* img, uint32_t lod, uint32_t row_begin, uint32_t row_end, void* peer_image_base, void* stream);

/* Copies `rows` rows of `width_bytes` bytes from local device memory (row pitch src_pitch) to peer-mapped memory (row pitch
 * dst_pitch) with the copy engines — no SM is involved, so a transfer queued beside running kernels costs them nothing.  Used for
 * the final gather of the batched views (vkr_views.PeerGather): a rank's block of per-view results lands in every peer's
 * gather buffer.  Asynchronous on `stream`. */
int vkr_peer_push_2d(const void* src, uint64_t src_pitch, void* peer_dst, uint64_t dst_pitch, uint64_t width_bytes, uint64_t rows, void* stream);

/* After everything queued on `stream` so far: system-scope release store of `value` to *flag (usually peer memory). */
int vkr_peer_signal(uint32_t* flag, uint32_t value, void* stream);
/* Holds `stream` until *flag >= value (system-scope acquire loads from a one-thread kernel).  After timeout_ms without
 * progress the kernel gives up and raises a sticky failure bit readable with vkr_peer_wait_failed() — it never hangs. */
int vkr_peer_wait(const uint32_t* flag, uint32_t value, uint32_t timeout_ms, void* stream);
/* Number of waits that timed out on the current device since the last call (resets the counter). Synchronises the device. */
int vkr_peer_wait_failed(void);

/* Epoch protocol of the fused gather (VkrLaunchExt.peer_delta in vkr_passes.h: the passes store their results into every rank's
 * arrays themselves).  Both kernels read the epoch from device memory, so a frame containing them replays as a CUDA graph.
 *   ctrl          one 32-bit word in THIS rank's memory: the number of epochs this rank has finished (0 at start)
 *   flags[r]      32-bit words in this rank's peer-visible memory: the last epoch rank r finished writing into this rank
 * vkr_peer_epoch_begin — before the frame's writers: waits (stream-side, with the timeout of vkr_peer_wait) until every other
 *   rank r < count has flags[r] >= ctrl[0], i.e. has finished with the buffer epoch ctrl[0] + 1 is about to overwrite.
 * vkr_peer_epoch_end — after them: ctrl[0] += 1 and that value is released (system scope) into remote_flags[r] for r < count
 *   (HOST array of peer-mapped addresses of this rank's flag word on rank r; NULL entries, e.g. the own slot, are skipped).
 * Consumer rule: read epoch k's gathered arrays on a stream ordered before this rank's epoch_end of k + 1.  At most 8 ranks. */
int vkr_peer_epoch_begin(const uint32_t* ctrl, const uint32_t* flags, uint32_t count, uint32_t self, uint32_t timeout_ms, void* stream);
int vkr_peer_epoch_end(uint32_t* ctrl, uint32_t* const* remote_flags, uint32_t count, void* stream);

#ifdef __cplusplus
}
#endif
#endif
