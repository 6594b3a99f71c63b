// rendergraph.cpp — see rendergraph.hpp.
#include "rendergraph.hpp"

#include <cuda_runtime.h>

#include <algorithm>

namespace rendergraph {

static void cuda_check(cudaError_t e, const char* what) {
  if (e != cudaSuccess) throw std::runtime_error(std::string(what) + ": " + cudaGetErrorString(e));
}

void vkr_check(int status, const char* what) {
  if (status != 0) throw std::runtime_error(std::string(what) + ": " + vkr_last_error());
}

RenderGraph::RenderGraph(int device, void* external_stream) : device_{device} {
  cuda_check(cudaSetDevice(device_), "cudaSetDevice");
  if (external_stream) {
    stream_ = external_stream;
    owns_stream_ = false;
  } else {
    cudaStream_t s;
    cuda_check(cudaStreamCreateWithFlags(&s, cudaStreamNonBlocking), "cudaStreamCreate");
    stream_ = s;
  }
}

RenderGraph::~RenderGraph() {
  cudaSetDevice(device_);
  if (stream_) {
    cudaStreamSynchronize((cudaStream_t)stream_);
    if (owns_stream_) cudaStreamDestroy((cudaStream_t)stream_);
  }
  for (void* p : allocations) cudaFree(p);
  for (auto& p : pending_) { cudaEventDestroy((cudaEvent_t)p.start); cudaEventDestroy((cudaEvent_t)p.stop); }
  for (void* e : event_pool_) cudaEventDestroy((cudaEvent_t)e);
  for (void* e : sync_events_) cudaEventDestroy((cudaEvent_t)e);
  if (aux_stream_) cudaStreamDestroy((cudaStream_t)aux_stream_);
  if (graph_exec_) cudaGraphExecDestroy((cudaGraphExec_t)graph_exec_);
}

void* RenderGraph::get_event() {
  if (!event_pool_.empty()) { void* e = event_pool_.back(); event_pool_.pop_back(); return e; }
  cudaEvent_t e;
  cuda_check(cudaEventCreate(&e), "cudaEventCreate");
  return e;
}

void RenderGraph::rebind_image(ImageResourceId id, const VkrImage& external) {
  const VkrImage& cur = image(id);
  if (!external.base || external.width != cur.width || external.height != cur.height || external.format != cur.format || external.mip_count != cur.mip_count)
    throw std::runtime_error("rebind_image: external image does not match the slot's descriptor");
  if (vkr_image_validate(&external) != VKR_OK) throw std::runtime_error(std::string("rebind_image: ") + vkr_last_error());
  slots[id.get_index()].desc = external;
}

const std::vector<RenderGraph::TaskTime>& RenderGraph::collect_task_times() {
  wait_idle();
  for (auto& p : pending_) {
    float ms = 0.f;
    cuda_check(cudaEventElapsedTime(&ms, (cudaEvent_t)p.start, (cudaEvent_t)p.stop), "cudaEventElapsedTime");
    auto it = std::find_if(totals_.begin(), totals_.end(), [&](const TaskTime& t) { return t.name == p.name; });
    if (it == totals_.end()) { totals_.push_back(TaskTime{p.name, 0.0, 0}); it = totals_.end() - 1; }
    it->ms += ms; it->count += 1;
    event_pool_.push_back(p.start); event_pool_.push_back(p.stop);
  }
  pending_.clear();
  return totals_;
}

void RenderGraph::reset_task_times() { collect_task_times(); totals_.clear(); }

ImageResourceId RenderGraph::create_image(const gpu::ImageInfo& info) {
  Slot s{};
  s.bytes = vkr_image_layout(info.width, info.height, info.format, info.mip_levels, &s.desc);
  if (s.bytes == 0) throw std::runtime_error("create_image: unsupported format / extent / mip count");
  s.layers = info.array_layers ? info.array_layers : 1;
  s.layer_stride = (s.bytes + 255) & ~(uint64_t)255;
  const uint64_t total = s.layers > 1 ? s.layer_stride * s.layers : s.bytes;
  cuda_check(cudaSetDevice(device_), "cudaSetDevice");
  cuda_check(cudaMalloc(&s.desc.base, total), "cudaMalloc(image)");
  cuda_check(cudaMemsetAsync(s.desc.base, 0, total, (cudaStream_t)stream_), "cudaMemset(image)");
  slots.push_back(s);
  allocations.push_back(s.desc.base);
  ImageResourceId id;
  id.index = (uint32_t)slots.size() - 1;
  return id;
}

BufferResourceId RenderGraph::create_buffer(uint64_t size) {
  if (size == 0) throw std::runtime_error("create_buffer: empty buffer");
  BufferSlot b{nullptr, size};
  cuda_check(cudaSetDevice(device_), "cudaSetDevice");
  cuda_check(cudaMalloc(&b.ptr, size), "cudaMalloc(buffer)");
  cuda_check(cudaMemsetAsync(b.ptr, 0, size, (cudaStream_t)stream_), "cudaMemset(buffer)");
  buffers.push_back(b);
  allocations.push_back(b.ptr);
  BufferResourceId id;
  id.index = (uint32_t)buffers.size() - 1;
  return id;
}

void* RenderGraph::buffer(BufferResourceId id) const {
  if (id.get_index() >= buffers.size()) throw std::runtime_error("Invalid buffer resource id");
  return buffers[id.get_index()].ptr;
}

uint64_t RenderGraph::buffer_size(BufferResourceId id) const {
  buffer(id);
  return buffers[id.get_index()].bytes;
}

const VkrImage& RenderGraph::image(ImageResourceId id) const {
  if (id.get_index() >= slots.size()) throw std::runtime_error("Invalid image resource id");
  return slots[id.get_index()].desc;
}

gpu::ImageInfo RenderGraph::get_descriptor(ImageResourceId id) const {
  const VkrImage& d = image(id);
  gpu::ImageInfo info;
  info.format = d.format; info.width = d.width; info.height = d.height; info.mip_levels = d.mip_count;
  info.array_layers = slots[id.get_index()].layers;
  return info;
}

uint64_t RenderGraph::layer_stride(ImageResourceId id) const {
  image(id);
  return slots[id.get_index()].layer_stride;
}

void RenderGraph::remap(ImageResourceId src, ImageResourceId dst) {
  image(src); image(dst);
  std::swap(slots[src.get_index()], slots[dst.get_index()]);
}

void RenderGraph::submit() {
  cuda_check(cudaSetDevice(device_), "cudaSetDevice");
  RenderResources resources{*this};
  last_names.clear();
  last_lanes.clear();
  std::vector<std::unique_ptr<BaseTask>> run;
  run.swap(tasks);  // a task may not enqueue tasks for this submit
  if (timing_ || !concurrent_ || run.size() < 3) {
    gpu::CmdContext cmd{stream_};
    for (auto& t : run) {
      last_names.push_back(t->name);
      last_lanes.push_back(0);
      if (timing_) {
        Pending p{t->name, get_event(), get_event()};
        cuda_check(cudaEventRecord((cudaEvent_t)p.start, (cudaStream_t)stream_), "cudaEventRecord");
        t->write_commands(resources, cmd);
        cuda_check(cudaEventRecord((cudaEvent_t)p.stop, (cudaStream_t)stream_), "cudaEventRecord");
        pending_.push_back(p);
      } else {
        t->write_commands(resources, cmd);
      }
    }
    fg_stats_.eager_submits++;
    return;
  }

  bool capture = frame_graph_ && warmed_;
  for (auto& t : run) capture = capture && t->capturable;
  if (capture) {
    cudaStream_t st = (cudaStream_t)stream_;
    if (cudaStreamBeginCapture(st, cudaStreamCaptureModeRelaxed) == cudaSuccess) {
      cudaGraph_t g = nullptr;
      bool ok = true;
      try {
        issue(run, resources);
      } catch (...) {
        cudaStreamEndCapture(st, &g);
        if (g) cudaGraphDestroy(g);
        cudaGetLastError();
        throw;
      }
      ok = cudaStreamEndCapture(st, &g) == cudaSuccess && g != nullptr;
      if (ok) {
        cudaGraphExec_t exec = (cudaGraphExec_t)graph_exec_;
        if (exec) {
          cudaGraphExecUpdateResultInfo info;
          if (cudaGraphExecUpdate(exec, g, &info) != cudaSuccess) {          // topology changed (another task list): start over
            cudaGetLastError();
            cudaGraphExecDestroy(exec);
            exec = nullptr;
          }
        }
        if (!exec) {
          ok = cudaGraphInstantiate(&exec, g, 0) == cudaSuccess;
          if (ok) fg_stats_.instantiations++;
        }
        graph_exec_ = ok ? exec : nullptr;
        cudaGraphDestroy(g);
        if (ok) ok = cudaGraphLaunch(exec, st) == cudaSuccess;
      }
      if (ok) { fg_stats_.launches++; return; }
      cudaGetLastError();
    }
    // capture / instantiate / launch refused: nothing of this frame reached the GPU; issue it eagerly and stop trying
    fg_stats_.capture_failures++;
    frame_graph_ = false;
    last_names.clear();
    last_lanes.clear();
  }
  issue(run, resources);
  warmed_ = true;                       // lazily built tables, module loads and the second stream exist from now on
  fg_stats_.eager_submits++;
}

void RenderGraph::issue(std::vector<std::unique_ptr<BaseTask>>& run, RenderResources& resources) {
  // Two lanes.  Dependencies at whole-image granularity from the declared uses: a read depends on the image's last
  // writer; a write on its last writer and on every reader since.
  if (!aux_stream_) {
    cudaStream_t s;
    cuda_check(cudaStreamCreateWithFlags(&s, cudaStreamNonBlocking), "cudaStreamCreate");
    aux_stream_ = s;
  }
  while (sync_events_.size() < run.size() + 1) {
    cudaEvent_t e;
    cuda_check(cudaEventCreateWithFlags(&e, cudaEventDisableTiming), "cudaEventCreate");
    sync_events_.push_back(e);
  }
  const int n = (int)run.size();
  // Hazards are keyed on the MEMORY a use touches, not on the slot index: two slots rebound to the same external allocation,
  // or layer views of one array image, are one resource.
  std::vector<const void*> res_keys;
  auto res_index = [&](uint32_t slot) -> size_t {
    const void* key = slot >= kBufferSlotBase ? buffers[slot - kBufferSlotBase].ptr : slots[slot].desc.base;
    for (size_t i = 0; i < res_keys.size(); i++) if (res_keys[i] == key) return i;
    res_keys.push_back(key);
    return res_keys.size() - 1;
  };
  const size_t n_res = slots.size() + buffers.size();
  std::vector<int> last_writer(n_res, -1), lane(n, 0);
  std::vector<std::vector<int>> readers(n_res);
  int last_barrier = -1;                 // last task that declared nothing: ordered against everything before AND after it
  void* lane_stream[2] = {stream_, aux_stream_};
  int lane_last[2] = {-1, -1};
  bool aux_used = false;
  // whatever happens below (a task may throw), the second stream is joined back into the graph's stream before we leave
  struct Join {
    RenderGraph* g; bool* used; int* last; 
    ~Join() { if (*used && *last >= 0) { cudaStreamWaitEvent((cudaStream_t)g->stream_, (cudaEvent_t)g->sync_events_[*last], 0); } }
  } join{this, &aux_used, &lane_last[1]};
  // the aux lane starts after everything already queued on the graph's stream
  cuda_check(cudaEventRecord((cudaEvent_t)sync_events_[n], (cudaStream_t)stream_), "cudaEventRecord");
  cuda_check(cudaStreamWaitEvent((cudaStream_t)aux_stream_, (cudaEvent_t)sync_events_[n], 0), "cudaStreamWaitEvent");
  for (int i = 0; i < n; i++) {
    BaseTask& t = *run[i];
    std::vector<int> deps;
    auto add = [&](int d) { if (d >= 0 && std::find(deps.begin(), deps.end(), d) == deps.end()) deps.push_back(d); };
    for (const ImageUse& u : t.uses) {
      add(last_writer[res_index(u.slot)]);
      if (u.write) for (int r : readers[res_index(u.slot)]) add(r);
    }
    auto is_dep = [&](int d) { return d >= 0 && std::find(deps.begin(), deps.end(), d) != deps.end(); };
    int l = 0;
    if (!t.uses.empty() && i > 0) {          // a task that declares nothing is ordered after everything: lane 0, all deps
      if (is_dep(lane_last[0]) || lane_last[0] < 0) l = 0;
      else if (is_dep(lane_last[1]) || lane_last[1] < 0) l = 1;
      else l = 1 - lane[i - 1];
    }
    if (t.uses.empty()) for (int d = 0; d < i; d++) add(d);
    add(last_barrier);
    lane[i] = l;
    for (int d : deps)
      if (lane[d] != l) cuda_check(cudaStreamWaitEvent((cudaStream_t)lane_stream[l], (cudaEvent_t)sync_events_[d], 0), "cudaStreamWaitEvent");
    gpu::CmdContext cmd{lane_stream[l]};
    t.write_commands(resources, cmd);
    cuda_check(cudaEventRecord((cudaEvent_t)sync_events_[i], (cudaStream_t)lane_stream[l]), "cudaEventRecord");
    lane_last[l] = i;
    aux_used |= l == 1;
    if (t.uses.empty()) last_barrier = i;
    for (const ImageUse& u : t.uses) {
      if (u.write) { last_writer[res_index(u.slot)] = i; readers[res_index(u.slot)].clear(); }
      else readers[res_index(u.slot)].push_back(i);
    }
    last_names.push_back(t.name);
    last_lanes.push_back(l);
  }
}

void RenderGraph::wait_idle() { cuda_check(cudaStreamSynchronize((cudaStream_t)stream_), "cudaStreamSynchronize"); }

ImageViewId RenderGraphBuilder::make_view(ImageResourceId id, uint32_t base_mip, uint32_t mip_count, bool write) {
  const VkrImage& d = graph.image(id);
  if (uses_) uses_->push_back(ImageUse{(uint32_t)id.get_index(), write});
  if (base_mip + mip_count > d.mip_count || mip_count == 0) throw std::runtime_error("Image subresource out of range");  // rendergraph.cpp:390-392
  gpu::ImageViewRange r;
  r.base_mip = base_mip; r.mips_count = mip_count;
  return ImageViewId{id, r};
}

ImageViewId RenderGraphBuilder::use_color_attachment(ImageResourceId id, uint32_t mip, uint32_t) { return make_view(id, mip, 1, true); }
ImageViewId RenderGraphBuilder::use_depth_attachment(ImageResourceId id, uint32_t mip, uint32_t) { return make_view(id, mip, 1, true); }
ImageViewId RenderGraphBuilder::use_storage_image(ImageResourceId id, uint32_t mip, uint32_t) { return make_view(id, mip, 1, true); }
ImageViewId RenderGraphBuilder::sample_image(ImageResourceId id, uint32_t base_mip, uint32_t mip_count, uint32_t, uint32_t) {
  return make_view(id, base_mip, mip_count, false);
}
ImageViewId RenderGraphBuilder::sample_image(ImageResourceId id) { return make_view(id, 0, graph.image(id).mip_count, false); }
gpu::ImageInfo RenderGraphBuilder::get_image_info(ImageResourceId id) { return graph.get_descriptor(id); }

void RenderGraphBuilder::use_buffer(BufferResourceId id, bool write) {
  graph.buffer(id);
  if (uses_) uses_->push_back(ImageUse{kBufferSlotBase + id.get_index(), write});
}
void RenderGraphBuilder::use_storage_buffer(BufferResourceId id, bool readonly) { use_buffer(id, !readonly); }
void RenderGraphBuilder::use_indirect_buffer(BufferResourceId id) { use_buffer(id, false); }
void RenderGraphBuilder::transfer_write(BufferResourceId id) { use_buffer(id, true); }

void RenderGraphBuilder::transfer_read(ImageResourceId id, uint32_t base_mip, uint32_t mip_count, uint32_t, uint32_t) { make_view(id, base_mip, mip_count, false); }
void RenderGraphBuilder::transfer_write(ImageResourceId id, uint32_t base_mip, uint32_t mip_count, uint32_t, uint32_t) { make_view(id, base_mip, mip_count, true); }

void* RenderResources::get_buffer(BufferResourceId id) const { return graph.buffer(id); }
const VkrImage* RenderResources::get_image(const ImageViewId& ref) const { return &graph.image(ref.get_id()); }
const VkrImage* RenderResources::get_image(ImageResourceId id) const { return &graph.image(id); }

}  // namespace rendergraph
