// launch_gap.cu — what does a kernel boundary cost on this box?
// The chain is ~10 dependent kernels per 1.2 ms frame; round-2 scaling runs showed identical kernel times but 0.12 ms more per
// frame on the 8-GPU box than on the 1- and 2-GPU boxes, at every N.  This measures, on ONE GPU of whatever box it runs on:
//   (1) N dependent empty kernels on one stream: time per kernel (launch throughput when the host runs ahead);
//   (2) the same N kernels as one CUDA graph: time per node;
//   (3) N dependent ~100 us spin kernels, stream and graph: (total - N * kernel time) / N = the gap between dependent kernels;
//   (4) fork/join: two streams joined by events per pair of kernels (the rendergraph's two lanes), stream and graph.
// nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/ubench/bin/launch_gap tools/ubench/launch_gap.cu
#include <cuda_runtime.h>
#include <stdio.h>
#include <stdlib.h>
#include <unistd.h>

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { printf("CUDA error %s at %s:%d\n", cudaGetErrorString(e_), __FILE__, __LINE__); exit(1); } } while (0)

__global__ void empty_kernel() {}
// streams `n` float4 through the SMs: read + write, ~2 * 16 * n bytes of HBM traffic (a stand-in for the chain's kernels)
__global__ void __launch_bounds__(256) copy_kernel(const float4* __restrict__ a, float4* __restrict__ b, size_t n) {
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) { float4 v = a[i]; v.x += 1.0f; b[i] = v; }
}
__global__ void spin_kernel(long long cycles, int* sink) {
  const long long t0 = clock64();
  while (clock64() - t0 < cycles) {}
  if (cycles < 0) *sink = 1;
}

static float elapsed(cudaEvent_t a, cudaEvent_t b) { float ms; CK(cudaEventElapsedTime(&ms, a, b)); return ms; }

int main(int argc, char** argv) {
  cudaDeviceProp prop;
  CK(cudaGetDeviceProperties(&prop, 0));
  int ndev = 0;
  CK(cudaGetDeviceCount(&ndev));
  printf("device %s, %d device(s) visible, %d SMs\n", prop.name, ndev, prop.multiProcessorCount);
  cudaStream_t s, s2;
  CK(cudaStreamCreate(&s)); CK(cudaStreamCreate(&s2));
  cudaEvent_t a, b;
  CK(cudaEventCreate(&a)); CK(cudaEventCreate(&b));
  int* sink; CK(cudaMalloc(&sink, 4));
  const int N = 1000;
  const dim3 grid(prop.multiProcessorCount * 4), block(128);

  // optional argv[1] = repetitions (default 2; the last one is reported; with several copies of this program started together on
  // different GPUs a larger count keeps them overlapping) and argv[2] = a file to wait for (common start)
  const int reps = argc > 1 ? atoi(argv[1]) : 2;
  if (argc > 2) { while (access(argv[2], F_OK) != 0) usleep(1000); }
  float acc[6] = {0, 0, 0, 0, 0, 0};
  for (int rep = 0; rep < reps; rep++) {                    // the first pass is the cold one
    // (1) stream, empty
    CK(cudaEventRecord(a, s));
    for (int i = 0; i < N; i++) empty_kernel<<<grid, block, 0, s>>>();
    CK(cudaEventRecord(b, s)); CK(cudaEventSynchronize(b));
    const float t_stream = elapsed(a, b) * 1e3f / N;
    // (2) graph, empty
    cudaGraph_t g; cudaGraphExec_t ge;
    CK(cudaStreamBeginCapture(s, cudaStreamCaptureModeRelaxed));
    for (int i = 0; i < N; i++) empty_kernel<<<grid, block, 0, s>>>();
    CK(cudaStreamEndCapture(s, &g)); CK(cudaGraphInstantiate(&ge, g, 0));
    CK(cudaGraphLaunch(ge, s)); CK(cudaStreamSynchronize(s));
    CK(cudaEventRecord(a, s)); CK(cudaGraphLaunch(ge, s)); CK(cudaEventRecord(b, s)); CK(cudaEventSynchronize(b));
    const float t_graph = elapsed(a, b) * 1e3f / N;
    CK(cudaGraphExecDestroy(ge)); CK(cudaGraphDestroy(g));
    // (3) ~100 us kernels
    const long long cyc = (long long)(100e-6 * prop.clockRate * 1e3);
    const int M = 100;
    CK(cudaEventRecord(a, s)); spin_kernel<<<grid, block, 0, s>>>(cyc, sink); CK(cudaEventRecord(b, s)); CK(cudaEventSynchronize(b));
    CK(cudaEventRecord(a, s)); spin_kernel<<<grid, block, 0, s>>>(cyc, sink); CK(cudaEventRecord(b, s)); CK(cudaEventSynchronize(b));
    const float one = elapsed(a, b) * 1e3f;
    CK(cudaEventRecord(a, s));
    for (int i = 0; i < M; i++) spin_kernel<<<grid, block, 0, s>>>(cyc, sink);
    CK(cudaEventRecord(b, s)); CK(cudaEventSynchronize(b));
    const float gap_stream = (elapsed(a, b) * 1e3f - M * one) / M;
    CK(cudaStreamBeginCapture(s, cudaStreamCaptureModeRelaxed));
    for (int i = 0; i < M; i++) spin_kernel<<<grid, block, 0, s>>>(cyc, sink);
    CK(cudaStreamEndCapture(s, &g)); CK(cudaGraphInstantiate(&ge, g, 0));
    CK(cudaGraphLaunch(ge, s)); CK(cudaStreamSynchronize(s));
    CK(cudaEventRecord(a, s)); CK(cudaGraphLaunch(ge, s)); CK(cudaEventRecord(b, s)); CK(cudaEventSynchronize(b));
    const float gap_graph = (elapsed(a, b) * 1e3f - M * one) / M;
    CK(cudaGraphExecDestroy(ge)); CK(cudaGraphDestroy(g));
    // (4) fork / join per pair: s: K ; s2 waits, K ; s waits s2 — the two-lane pattern, M/2 times
    cudaEvent_t ev[2];
    CK(cudaEventCreateWithFlags(&ev[0], cudaEventDisableTiming)); CK(cudaEventCreateWithFlags(&ev[1], cudaEventDisableTiming));
    auto forkjoin = [&]() {
      for (int i = 0; i < M / 2; i++) {
        spin_kernel<<<grid, block, 0, s>>>(cyc, sink);
        CK(cudaEventRecord(ev[0], s)); CK(cudaStreamWaitEvent(s2, ev[0], 0));
        spin_kernel<<<grid, block, 0, s2>>>(cyc, sink);
        CK(cudaEventRecord(ev[1], s2)); CK(cudaStreamWaitEvent(s, ev[1], 0));
      }
    };
    CK(cudaEventRecord(a, s)); forkjoin(); CK(cudaEventRecord(b, s)); CK(cudaEventSynchronize(b));
    const float gap_fj_stream = (elapsed(a, b) * 1e3f - M * one) / M;
    CK(cudaStreamBeginCapture(s, cudaStreamCaptureModeRelaxed));
    forkjoin();
    CK(cudaStreamEndCapture(s, &g)); CK(cudaGraphInstantiate(&ge, g, 0));
    CK(cudaGraphLaunch(ge, s)); CK(cudaStreamSynchronize(s));
    CK(cudaEventRecord(a, s)); CK(cudaGraphLaunch(ge, s)); CK(cudaEventRecord(b, s)); CK(cudaEventSynchronize(b));
    const float gap_fj_graph = (elapsed(a, b) * 1e3f - M * one) / M;
    CK(cudaGraphExecDestroy(ge)); CK(cudaGraphDestroy(g));
    // (5) memory-streaming kernels (256 MB read + 256 MB written each, ~85 us): 100 dependent launches, ping-pong buffers
    {
      static float4 *bufA = nullptr, *bufB = nullptr;
      const size_t n = (size_t)16 << 20;
      if (!bufA) { CK(cudaMalloc(&bufA, n * 16)); CK(cudaMalloc(&bufB, n * 16)); CK(cudaMemset(bufA, 0, n * 16)); }
      const dim3 cg(prop.multiProcessorCount * 8);
      CK(cudaEventRecord(a, s)); copy_kernel<<<cg, 256, 0, s>>>(bufA, bufB, n); CK(cudaEventRecord(b, s)); CK(cudaEventSynchronize(b));
      CK(cudaEventRecord(a, s)); copy_kernel<<<cg, 256, 0, s>>>(bufA, bufB, n); CK(cudaEventRecord(b, s)); CK(cudaEventSynchronize(b));
      const float onec = elapsed(a, b) * 1e3f;
      CK(cudaEventRecord(a, s));
      for (int i = 0; i < M; i++) copy_kernel<<<cg, 256, 0, s>>>((i & 1) ? bufB : bufA, (i & 1) ? bufA : bufB, n);
      CK(cudaEventRecord(b, s)); CK(cudaEventSynchronize(b));
      const float tot = elapsed(a, b) * 1e3f;
      if (rep == reps - 1) printf("memory-streaming kernel: %.1f us alone, %.1f us each in a chain of %d dependent launches (extra %.2f us per launch)\n", onec, tot / M, M, tot / M - onec);
    }
    if (rep >= 1) { acc[0] += t_stream; acc[1] += t_graph; acc[2] += gap_stream; acc[3] += gap_graph; acc[4] += gap_fj_stream; acc[5] += gap_fj_graph; }
    if (rep == reps - 1) {
      const float n = (float)(reps - 1);
      printf("mean of %d passes | empty kernel: %.2f us each in a stream, %.2f us each in a graph | extra time per dependent ~%.0f us kernel: stream %.2f us, graph %.2f us | "
             "across two streams (event fork/join): stream %.2f us, graph %.2f us\n", reps - 1, acc[0] / n, acc[1] / n, one, acc[2] / n, acc[3] / n, acc[4] / n, acc[5] / n);
    }
  }
  return 0;
}
