// A task that declares no uses is a barrier: ordered after everything before it AND before everything after it
// (it used to be one-sided — a later task placed on the second stream waited only for its declared producers).
//   A  declares write X            : X = 1
//   B  declares nothing            : scrubs a large buffer (slow), then X = 2
//   C  declares read X, write Y    : Y = X
// With the one-sided barrier C went to the second stream waiting for A only and could copy X before B wrote it.
// Also exercises hazards keyed on memory: D reads Y through a SECOND slot rebound to Y's allocation and must see C's write.
#include <cstdio>
#include <cstring>
#include <vector>
#include <cuda_runtime.h>
#include "rendergraph.hpp"

using namespace rendergraph;
struct Nothing {};

int main() {
  RenderGraph graph(0);
  gpu::ImageInfo info; info.format = VKR_FORMAT_R32_SFLOAT; info.width = 1024; info.height = 1024; info.mip_levels = 1; info.array_layers = 1;
  ImageResourceId X = graph.create_image(info), Y = graph.create_image(info), Yalias = graph.create_image(info), Z = graph.create_image(info);
  graph.rebind_image(Yalias, graph.image(Y));               // two slots, one allocation
  void* big = nullptr;
  const size_t big_bytes = (size_t)1 << 30;
  if (cudaMalloc(&big, big_bytes) != cudaSuccess) { printf("alloc failed\n"); return 2; }
  const size_t bytes = (size_t)graph.image(X).row_pitch[0] * 1024;
  int bad = 0;
  for (int frame = 0; frame < 6; frame++) {                  // frame 0 eager, later frames through the frame graph
    graph.add_task<Nothing>("A", [&](Nothing&, RenderGraphBuilder& b) { b.use_storage_image(X, 0, 0); },
                            [=, &graph](Nothing&, RenderResources& r, gpu::CmdContext& c) { cudaMemsetAsync(r.get_image(X)->base, 1, bytes, (cudaStream_t)c.stream); });
    graph.add_task<Nothing>("B", [&](Nothing&, RenderGraphBuilder&) {},
                            [=, &graph](Nothing&, RenderResources& r, gpu::CmdContext& c) {
                              cudaMemsetAsync(big, frame, big_bytes, (cudaStream_t)c.stream);
                              cudaMemsetAsync(r.get_image(X)->base, 2, bytes, (cudaStream_t)c.stream);
                            });
    graph.add_task<Nothing>("C", [&](Nothing&, RenderGraphBuilder& b) { b.sample_image(X); b.use_storage_image(Y, 0, 0); },
                            [=, &graph](Nothing&, RenderResources& r, gpu::CmdContext& c) {
                              cudaMemcpyAsync(r.get_image(Y)->base, r.get_image(X)->base, bytes, cudaMemcpyDeviceToDevice, (cudaStream_t)c.stream);
                            });
    graph.add_task<Nothing>("D", [&](Nothing&, RenderGraphBuilder& b) { b.sample_image(Yalias); b.use_storage_image(Z, 0, 0); },
                            [=, &graph](Nothing&, RenderResources& r, gpu::CmdContext& c) {
                              cudaMemcpyAsync(r.get_image(Z)->base, r.get_image(Yalias)->base, bytes, cudaMemcpyDeviceToDevice, (cudaStream_t)c.stream);
                            });
    graph.submit();
    graph.wait_idle();
    std::vector<unsigned char> h(bytes);
    cudaMemcpy(h.data(), graph.image(Z).base, bytes, cudaMemcpyDeviceToHost);
    size_t wrong = 0;
    for (unsigned char v : h) wrong += v != 2;
    if (wrong) { printf("frame %d: %zu of %zu bytes of Z are not 2 (first %d)\n", frame, wrong, bytes, (int)h[0]); bad++; }
    cudaMemset(graph.image(Z).base, 0, bytes);
    cudaMemset(graph.image(Y).base, 0, bytes);
  }
  const auto& st = graph.frame_graph_stats();
  printf("lanes of the last frame:");
  for (int l : graph.last_task_lanes()) printf(" %d", l);
  printf(" | graph launches %llu, eager %llu, capture failures %llu\n", (unsigned long long)st.launches, (unsigned long long)st.eager_submits, (unsigned long long)st.capture_failures);
  cudaFree(big);
  printf(bad ? "FAILED\n" : "ok\n");
  return bad ? 1 : 0;
}
