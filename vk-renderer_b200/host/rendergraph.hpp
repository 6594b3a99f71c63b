// rendergraph.hpp — headless re-creation of the interface the reference passes bind through
// (reference src/rendergraph/rendergraph.hpp:17-158, src/rendergraph/resources.hpp:15-118,
// src/gpu/resource_info.hpp:11-45).  Same shape — create_image / get_descriptor / add_task<T>(name,
// setup_cb, exec_cb) / submit / remap, ImageResourceId slots that must be re-resolved every frame —
// but the Vulkan machinery behind it (barrier derivation, VkEvents, descriptor pools, swapchain) is
// replaced by CUDA stream order: submit() replays the task list in FIFO order onto one stream.
// Images are linear device buffers (VkrImage); the graph owns them, passes hold only ids.
#pragma once
#include <cstdint>
#include <functional>
#include <memory>
#include <stdexcept>
#include <string>
#include <vector>

#include "../../include/vkr_passes.h"

namespace gpu {

// reference src/gpu/resource_info.hpp:11-45 (type / aspect dropped: every view here is 2D)
struct ImageViewRange {
  uint32_t base_mip = 0;
  uint32_t mips_count = 1;
  uint32_t base_layer = 0;
  uint32_t layers_count = 1;
};

// reference gpu::ImageInfo as the passes use it: {format, aspect, width, height, depth, mips, layers}
struct ImageInfo {
  uint32_t format = 0;  // VkrFormat (= VkFormat value)
  uint32_t width = 0, height = 0;
  uint32_t mip_levels = 1;
  uint32_t array_layers = 1;
  struct Extent2D { uint32_t width, height; };
  Extent2D extent2D() const { return {width, height}; }
};

// Stands in for gpu::CmdContext (reference src/gpu/cmd_buffers.hpp:162-247): the only thing a task
// body needs to issue work is the stream of the frame being recorded.
struct CmdContext {
  void* stream = nullptr;  // cudaStream_t
};

}  // namespace gpu

namespace rendergraph {

struct GraphResources;

struct ImageResourceId {
  ImageResourceId() {}
  uint32_t get_index() const { return index; }
  bool operator==(const ImageResourceId& id) const { return index == id.index; }

 private:
  uint32_t index = ~0u;
  friend struct GraphResources;
  friend struct RenderGraph;
};

struct ImageViewId {
  ImageViewId() {}
  ImageViewId(ImageResourceId id, gpu::ImageViewRange view) : res_id{id}, range{view} {}
  ImageResourceId get_id() const { return res_id; }
  const gpu::ImageViewRange& get_range() const { return range; }
  operator ImageResourceId() const { return res_id; }

 private:
  ImageResourceId res_id;
  gpu::ImageViewRange range;
};

// reference src/rendergraph/resources.hpp: a slot index like ImageResourceId, for linear device buffers (SSBOs, indirect
// dispatch commands).  Only the SSSR tile path uses buffers (reference src/advanced_ssr.cpp:77-83).
struct BufferResourceId {
  BufferResourceId() {}
  uint32_t get_index() const { return index; }
  bool operator==(const BufferResourceId& id) const { return index == id.index; }

 private:
  uint32_t index = ~0u;
  friend struct RenderGraph;
};

struct RenderGraph;

// One declared use of an image by a task: whole-image granularity (slot index at record time), read or write.
// Buffers share the numbering: slot = kBufferSlotBase + buffer index.
struct ImageUse { uint32_t slot; bool write; };
constexpr uint32_t kBufferSlotBase = 0x40000000u;

// reference rendergraph.hpp:17-55.  Usage declarations keep their names; they validate the subresource range
// (reference rendergraph.cpp:390-392 "Image subresource out of range") and record the image as read or written by the
// task.  Where the reference derives layout transitions and barriers from these declarations, submit() derives which
// tasks of a frame may run side by side on two CUDA streams (see RenderGraph::submit).
struct RenderGraphBuilder {
  explicit RenderGraphBuilder(RenderGraph& g, std::vector<ImageUse>* uses = nullptr) : graph{g}, uses_{uses} {}
  ImageViewId use_color_attachment(ImageResourceId id, uint32_t mip, uint32_t layer);
  ImageViewId use_depth_attachment(ImageResourceId id, uint32_t mip, uint32_t layer);
  ImageViewId use_storage_image(ImageResourceId id, uint32_t mip, uint32_t layer);
  ImageViewId sample_image(ImageResourceId id, uint32_t base_mip, uint32_t mip_count, uint32_t base_layer, uint32_t layer_count);
  ImageViewId sample_image(ImageResourceId id);  // whole image
  gpu::ImageInfo get_image_info(ImageResourceId id);
  // reference rendergraph.hpp:40-55 (shader stage / usage flags dropped): declare a buffer as read or written by the task
  void use_storage_buffer(BufferResourceId id, bool readonly = true);
  void use_indirect_buffer(BufferResourceId id);
  void transfer_write(BufferResourceId id);
  // image transfers (reference rendergraph.hpp: transfer_read / transfer_write with a subresource range)
  void transfer_read(ImageResourceId id, uint32_t base_mip, uint32_t mip_count, uint32_t base_layer, uint32_t layer_count);
  void transfer_write(ImageResourceId id, uint32_t base_mip, uint32_t mip_count, uint32_t base_layer, uint32_t layer_count);

 private:
  RenderGraph& graph;
  std::vector<ImageUse>* uses_;
  void use_buffer(BufferResourceId id, bool write);
  ImageViewId make_view(ImageResourceId id, uint32_t base_mip, uint32_t mip_count, bool write);
};

// reference rendergraph.hpp:57-83: resolves ids to memory at replay time (after any remap).
struct RenderResources {
  explicit RenderResources(RenderGraph& g) : graph{g} {}
  // The image behind a view, as the C ABI wants it (whole image; the pass passes the mip window on).
  const VkrImage* get_image(const ImageViewId& ref) const;
  const VkrImage* get_image(ImageResourceId id) const;
  void* get_buffer(BufferResourceId id) const;   // device pointer (reference: gpu::BufferPtr -> api_buffer())

 private:
  RenderGraph& graph;
};

struct BaseTask {
  explicit BaseTask(const std::string& task_name) : name{task_name} {}
  virtual void write_commands(RenderResources&, gpu::CmdContext&) = 0;
  virtual ~BaseTask() {}
  std::string name;
  std::vector<ImageUse> uses;   // filled by the builder while the task is declared
  bool capturable = true;       // false: records events the host waits on / touches host state -> the frame is issued eagerly
};

template <typename TaskData>
using TaskRunCB = std::function<void(TaskData&, RenderResources&, gpu::CmdContext&)>;
template <typename TaskData>
using TaskCreateCB = std::function<void(TaskData&, RenderGraphBuilder&)>;

template <typename TaskData>
struct Task : BaseTask {
  explicit Task(const std::string& name) : BaseTask{name} {}
  TaskData data;
  TaskRunCB<TaskData> callback;
  void write_commands(RenderResources& resources, gpu::CmdContext& cmd) override { callback(data, resources, cmd); }
};

struct RenderGraph {
  // external_stream: run on the caller's cudaStream_t instead of a stream owned by the graph
  explicit RenderGraph(int device = 0, void* external_stream = nullptr);
  ~RenderGraph();
  RenderGraph(const RenderGraph&) = delete;
  RenderGraph& operator=(const RenderGraph&) = delete;

  template <typename TaskData>
  void add_task(const std::string& name, TaskCreateCB<TaskData> create_cb, TaskRunCB<TaskData> run_cb) {  // reference :116-128
    std::unique_ptr<Task<TaskData>> ptr{new Task<TaskData>{name}};
    RenderGraphBuilder builder{*this, &ptr->uses};
    create_cb(ptr->data, builder);
    ptr->callback = run_cb;
    tasks.push_back(std::move(ptr));
  }

  // Replays the recorded tasks and clears the list (reference rendergraph.cpp:271-337 minus layout transitions /
  // present).  Asynchronous; wait_idle() joins.  Tasks are issued in recording order.  With concurrent branches enabled
  // (default) a task none of whose declared images was written by — or, for the images it writes, read by — the task
  // issued just before it goes to a second stream, so independent branches of the frame (here the SSR filter + blur
  // against the three GTAO tasks) share the GPU and fill each other's tails and latency-bound phases; every true
  // dependency across the two streams is carried by an event, and the second stream is joined back into the graph's
  // stream before submit() returns.  Per-task timing needs tasks that do not overlap: while it is on, or with
  // concurrent branches off, everything runs on the one stream.
  void submit();
  void enable_concurrent_branches(bool on) { concurrent_ = on; }
  // Frame graph (default on): from the second submit on, the two-lane issue sequence is stream-captured every frame and
  // applied to ONE instantiated CUDA graph with cudaGraphExecUpdate (same topology; kernel parameters, grids and the
  // ping-ponged image pointers change), then launched: the whole frame reaches the GPU as one unit, so the gaps between
  // its ~10 kernels no longer depend on how fast this thread enqueues them (reference analogue: one command buffer per
  // frame, src/rendergraph/rendergraph.cpp:271-337).  Frames that hold a task marked not capturable, per-task timing and
  // single-stream mode are issued eagerly as before; a refused update re-instantiates, a failed capture turns the
  // feature off and re-issues the frame eagerly (the chain's tasks are pure C-ABI calls).
  void enable_frame_graph(bool on) { frame_graph_ = on; }
  void mark_last_task_not_capturable() { if (!tasks.empty()) tasks.back()->capturable = false; }
  struct FrameGraphStats { uint64_t launches = 0, instantiations = 0, eager_submits = 0, capture_failures = 0; };
  const FrameGraphStats& frame_graph_stats() const { return fg_stats_; }
  // lane (0 = the graph's stream, 1 = the second one) each task of the last submit ran on
  const std::vector<int>& last_task_lanes() const { return last_lanes; }
  void wait_idle();

  ImageResourceId create_image(const gpu::ImageInfo& info);  // zero-initialised device memory
  gpu::ImageInfo get_descriptor(ImageResourceId id) const;
  // reference rendergraph.hpp create_buffer(VmaMemoryUsage, size, usage): zero-initialised device memory
  BufferResourceId create_buffer(uint64_t size);
  void* buffer(BufferResourceId id) const;
  uint64_t buffer_size(BufferResourceId id) const;
  // Array images (gpu::ImageInfo::array_layers, reference src/gpu/resource_info.hpp:11-19): layer l of slot `id` lies
  // layer_stride(id) * l bytes after the descriptor's base.
  uint64_t layer_stride(ImageResourceId id) const;
  void remap(ImageResourceId src, ImageResourceId dst);      // reference resources.cpp:89-91: swap slot contents

  // Point a slot at caller-owned device memory (same extent / format / mips): lets a resident input
  // set be bound without a copy.  The graph never frees borrowed memory; its own allocation stays alive.
  void rebind_image(ImageResourceId id, const VkrImage& external);

  // Per-task device timing: cudaEvents recorded around every task replayed by submit().
  void enable_task_timing(bool on) { timing_ = on; }
  struct TaskTime { std::string name; double ms = 0; uint64_t count = 0; };
  // Waits for the stream, folds the pending event pairs into the totals and returns them.
  const std::vector<TaskTime>& collect_task_times();
  void reset_task_times();

  const VkrImage& image(ImageResourceId id) const;
  void* stream() const { return stream_; }
  // reference RenderGraph::get_frames_count(): frames in flight (swapchain images there).  submit() here replays onto a stream
  // immediately, so one frame is in flight; ReadBackSystem waits get_frames_count() + 1 submits like the reference.
  uint32_t get_frames_count() const { return 1; }
  const std::vector<std::string>& last_task_names() const { return last_names; }

 private:
  struct Slot { VkrImage desc; uint64_t bytes; uint32_t layers = 1; uint64_t layer_stride = 0; };
  std::vector<Slot> slots;
  struct BufferSlot { void* ptr; uint64_t bytes; };
  std::vector<BufferSlot> buffers;
  std::vector<void*> allocations;   // everything create_image() allocated (slots may be rebound / swapped)
  bool timing_ = false;
  struct Pending { std::string name; void* start; void* stop; };
  std::vector<Pending> pending_;
  std::vector<void*> event_pool_;
  std::vector<TaskTime> totals_;
  void* get_event();
  std::vector<std::unique_ptr<BaseTask>> tasks;
  std::vector<std::string> last_names;
  std::vector<int> last_lanes;
  bool concurrent_ = true;
  void issue(std::vector<std::unique_ptr<BaseTask>>& run, RenderResources& resources);   // the two-lane issue sequence
  bool frame_graph_ = true, warmed_ = false;
  void* graph_exec_ = nullptr;          // cudaGraphExec_t, updated every captured frame
  FrameGraphStats fg_stats_;
  void* aux_stream_ = nullptr;          // created on first use
  std::vector<void*> sync_events_;      // cudaEventDisableTiming, one per task of a submit, reused
  void* stream_ = nullptr;
  bool owns_stream_ = true;
  int device_ = 0;
  friend struct RenderGraphBuilder;
  friend struct RenderResources;
};

// VKCHECK analogue (reference src/gpu/common.hpp:9): non-zero C-ABI status -> std::runtime_error
void vkr_check(int status, const char* what);

}  // namespace rendergraph
