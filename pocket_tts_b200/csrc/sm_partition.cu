// Spatial SM partition for the two branches of the frame loop (CUDA green contexts, driver API >= 12.4):
// the FlowLM chain (many short dependent kernels, latency critical) gets `flow_sms` SMs, the Mimi branch (wide
// throughput kernels) the rest, so that a resident Mimi grid can never make a FlowLM kernel wait for an SM.
#include <cuda.h>
#include <cuda_runtime.h>

#include <cstdio>

#include "engine.h"

namespace ptts {

namespace {
template <typename F>
bool entry(const char* name, F* fn) {
  void* p = nullptr;
  cudaDriverEntryPointQueryResult q;
  if (cudaGetDriverEntryPoint(name, &p, cudaEnableDefault, &q) != cudaSuccess || q != cudaDriverEntryPointSuccess || !p) return false;
  *fn = reinterpret_cast<F>(p);
  return true;
}
}  // namespace

// Returns false (and leaves the outputs untouched) when the driver cannot partition; the caller then keeps its
// ordinary streams.  On success both streams are non-blocking streams of two green contexts with disjoint SMs.
bool make_sm_partition(int device, int flow_sms, int prio_flow, int prio_mimi, cudaStream_t* s_flow, cudaStream_t* s_mimi,
                       int* n_flow, int* n_mimi) {
  typedef CUresult (*GetRes)(CUdevice, CUdevResource*, CUdevResourceType);
  typedef CUresult (*Split)(CUdevResource*, unsigned int*, const CUdevResource*, CUdevResource*, unsigned int, unsigned int);
  typedef CUresult (*GenDesc)(CUdevResourceDesc*, CUdevResource*, unsigned int);
  typedef CUresult (*GreenCreate)(CUgreenCtx*, CUdevResourceDesc, CUdevice, unsigned int);
  typedef CUresult (*GreenStream)(CUstream*, CUgreenCtx, unsigned int, int);
  typedef CUresult (*DevGet)(CUdevice*, int);
  GetRes get_res; Split split; GenDesc gen_desc; GreenCreate green_create; GreenStream green_stream; DevGet dev_get;
  if (!entry("cuDeviceGetDevResource", &get_res) || !entry("cuDevSmResourceSplitByCount", &split) ||
      !entry("cuDevResourceGenerateDesc", &gen_desc) || !entry("cuGreenCtxCreate", &green_create) ||
      !entry("cuGreenCtxStreamCreate", &green_stream) || !entry("cuDeviceGet", &dev_get))
    return false;
  cudaFree(0);   // primary context
  CUdevice dev;
  if (dev_get(&dev, device) != CUDA_SUCCESS) return false;
  CUdevResource all, flow, rest;
  if (get_res(dev, &all, CU_DEV_RESOURCE_TYPE_SM) != CUDA_SUCCESS) return false;
  unsigned int groups = 1;
  if (split(&flow, &groups, &all, &rest, 0, (unsigned)flow_sms) != CUDA_SUCCESS || groups != 1) return false;
  if (flow.sm.smCount < 8 || rest.sm.smCount < 8) return false;
  CUdevResourceDesc d_flow, d_rest;
  if (gen_desc(&d_flow, &flow, 1) != CUDA_SUCCESS || gen_desc(&d_rest, &rest, 1) != CUDA_SUCCESS) return false;
  CUgreenCtx g_flow, g_rest;   // live for the life of the process (a handful of engines at most)
  if (green_create(&g_flow, d_flow, dev, CU_GREEN_CTX_DEFAULT_STREAM) != CUDA_SUCCESS) return false;
  if (green_create(&g_rest, d_rest, dev, CU_GREEN_CTX_DEFAULT_STREAM) != CUDA_SUCCESS) return false;
  CUstream a, b;
  if (green_stream(&a, g_flow, CU_STREAM_NON_BLOCKING, prio_flow) != CUDA_SUCCESS) return false;
  if (green_stream(&b, g_rest, CU_STREAM_NON_BLOCKING, prio_mimi) != CUDA_SUCCESS) return false;
  *s_flow = a; *s_mimi = b;
  *n_flow = (int)flow.sm.smCount; *n_mimi = (int)rest.sm.smCount;
  return true;
}

}  // namespace ptts
