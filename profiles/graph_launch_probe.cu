// Developer probe: how fast does a B200 issue the kernel nodes of CUDA graphs made of short dependent kernels?
//   nvcc -gencode arch=compute_100a,code=sm_100a -O2 -o graph_launch_probe graph_launch_probe.cu
// Varies: nodes per graph, number of distinct executable graphs used in rotation, a small memcpy / an event
// record between launches, and a second stream running its own graphs.  Prints us per node.
#include <cuda_runtime.h>

#include <cstdio>
#include <vector>

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { printf("CUDA error %s at %d\n", cudaGetErrorString(e_), __LINE__); return 1; } } while (0)

__global__ void tiny(float* p, int work) {
  float v = p[threadIdx.x];
  for (int i = 0; i < work; ++i) v = v * 1.0001f + 0.5f;
  p[threadIdx.x] = v;
}

static cudaGraphExec_t make_graph(cudaStream_t st, float* buf, int nodes, int blocks, int work, bool pdl) {
  cudaGraph_t g;
  cudaStreamBeginCapture(st, cudaStreamCaptureModeThreadLocal);
  for (int i = 0; i < nodes; ++i) {
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = dim3(blocks); cfg.blockDim = dim3(128); cfg.stream = st;
    cudaLaunchAttribute a[1];
    a[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    a[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = a; cfg.numAttrs = pdl ? 1 : 0;
    cudaLaunchKernelEx(&cfg, tiny, buf, work);
  }
  cudaStreamEndCapture(st, &g);
  cudaGraphExec_t e;
  cudaGraphInstantiate(&e, g, 0);
  cudaGraphDestroy(g);
  return e;
}

int main() {
  cudaStream_t s1, s2;
  int lo, hi;
  CK(cudaDeviceGetStreamPriorityRange(&lo, &hi));
  CK(cudaStreamCreateWithPriority(&s1, cudaStreamNonBlocking, hi));
  CK(cudaStreamCreateWithPriority(&s2, cudaStreamNonBlocking, lo));
  float *buf, *buf2, *src, *dst;
  CK(cudaMalloc(&buf, 4096)); CK(cudaMalloc(&buf2, 4096)); CK(cudaMalloc(&src, 8192)); CK(cudaMalloc(&dst, 8192));
  CK(cudaMemset(buf, 0, 4096)); CK(cudaMemset(buf2, 0, 4096));
  cudaEvent_t e0, e1, ev;
  CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1)); CK(cudaEventCreateWithFlags(&ev, cudaEventDisableTiming));
  const int launches = 64;
  printf("%-6s %-6s %-7s %-7s %-7s %-8s  us/launch  us/node\n", "nodes", "execs", "memcpy", "event", "pdl", "2nd-strm");
  for (int nodes : {35, 71, 106, 142, 284})
    for (int execs : {1, 2, 8})
      for (int variant = 0; variant < 5; ++variant) {
        const bool memcpy_between = variant == 1 || variant == 3, event_between = variant == 2 || variant == 3;
        const bool second = variant == 4;
        if ((nodes == 35 || nodes == 284) && variant != 0) continue;
        for (int pdl = 0; pdl < 2; ++pdl) {
          if (pdl && variant != 0) continue;
          std::vector<cudaGraphExec_t> g(execs);
          for (auto& x : g) x = make_graph(s1, buf, nodes, 64, 200, pdl);
          cudaGraphExec_t g2 = second ? make_graph(s2, buf2, 35, 148, 2000, false) : nullptr;
          for (int rep = 0; rep < 2; ++rep) {
            CK(cudaStreamSynchronize(s1)); CK(cudaStreamSynchronize(s2));
            CK(cudaEventRecord(e0, s1));
            for (int i = 0; i < launches; ++i) {
              if (memcpy_between) CK(cudaMemcpyAsync(dst, src, 8192, cudaMemcpyDeviceToDevice, s1));
              CK(cudaGraphLaunch(g[i % execs], s1));
              if (event_between) CK(cudaEventRecord(ev, s1));
              if (second) CK(cudaGraphLaunch(g2, s2));
            }
            CK(cudaEventRecord(e1, s1));
            CK(cudaStreamSynchronize(s1)); CK(cudaStreamSynchronize(s2));
          }
          float ms = 0;
          CK(cudaEventElapsedTime(&ms, e0, e1));
          printf("%-6d %-6d %-7d %-7d %-7d %-8d  %8.1f  %7.2f\n", nodes, execs, (int)memcpy_between, (int)event_between, pdl,
                 (int)second, ms * 1e3 / launches, ms * 1e3 / launches / nodes);
          for (auto& x : g) cudaGraphExecDestroy(x);
          if (g2) cudaGraphExecDestroy(g2);
        }
      }
  return 0;
}
