// ptts_pcm_to_int16_static: the PCM egress conversion (data/audio.py:77-78 on the device, see ptts_pcm_to_int16) for
// callers that must not allocate — code under CUDA graph capture, tight serving loops, harnesses that forbid
// cudaMalloc; no counterpart in the reference.
//
// ptts_pcm_to_int16 stages HOST buffers through a device buffer that it grows with cudaMalloc / cudaFree.  Here the
// staging scratch is a file-scope device array of fixed size instead: every GPU that loads the library holds
// PTTS_PCM_STATIC_MAX_SAMPLES * 6 bytes (1.5 MiB) until the process exits.  This file calls no cudaMalloc, cudaFree,
// cudaMallocAsync or any other allocating routine.
//
//   * The array's name is a host-side stand-in in host code, not a device address: its device address is resolved
//     with cudaGetSymbolAddress once per device (with that device current) and cached.
//   * Every call on a GPU shares the one array and the array keeps the previous call's contents.  Nothing has to
//     start at zero: a call reads only the n samples it has just copied in and copies out only the n samples its
//     kernel has just written.  A call returns after its stream has drained, so calls on one engine never overlap;
//     calls that do overlap on the same GPU (two engines / two host threads) RACE on the array — serialise them.
#include <mutex>

#include "engine.h"

namespace ptts {

namespace {

constexpr long long kMaxSamples = PTTS_PCM_STATIC_MAX_SAMPLES;
constexpr size_t kInBytes = (size_t)kMaxSamples * sizeof(float);   // fp32 samples in, then int16 samples out
constexpr size_t kOutBytes = (size_t)kMaxSamples * sizeof(short);

static __device__ __align__(256) unsigned char g_pcm_scratch[kInBytes + kOutBytes];

constexpr int kMaxDevices = 64;
std::mutex g_mu;
unsigned char* g_scratch_addr[kMaxDevices] = {};   // device address of g_pcm_scratch, per CUDA device

// device address of the array on the CURRENT device (resolved once per device)
cudaError_t scratch_address(unsigned char** out) {
  int dev = -1;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return e;
  if (dev < 0 || dev >= kMaxDevices) return cudaErrorInvalidDevice;
  std::lock_guard<std::mutex> lock(g_mu);
  if (!g_scratch_addr[dev]) {
    void* p = nullptr;
    e = cudaGetSymbolAddress(&p, g_pcm_scratch);
    if (e != cudaSuccess) return e;
    g_scratch_addr[dev] = static_cast<unsigned char*>(p);
  }
  *out = g_scratch_addr[dev];
  return cudaSuccess;
}

bool on_device(const void* p) {
  cudaPointerAttributes a{};
  if (cudaPointerGetAttributes(&a, p) != cudaSuccess) { cudaGetLastError(); return false; }
  return a.type == cudaMemoryTypeDevice || a.type == cudaMemoryTypeManaged;
}

}  // namespace

// The engine's device is current (the C-ABI wrapper holds the device guard).
int pcm_to_int16_static(cudaStream_t stream, const float* pcm_any, short* out_any, long long n, long long* launches) {
  if (n < 0 || (n > 0 && (!pcm_any || !out_any))) PTTS_FAIL(PTTS_EINVAL, "pcm_to_int16_static: bad arguments");
  if (n > kMaxSamples)
    PTTS_FAIL(PTTS_EINVAL, "pcm_to_int16_static: %lld samples exceed the static scratch of %lld samples "
              "(PTTS_PCM_STATIC_MAX_SAMPLES); split the call or use ptts_pcm_to_int16", n, kMaxSamples);
  if (n == 0) return PTTS_OK;
  unsigned char* scratch = nullptr;
  PTTS_CUDA_CHECK(scratch_address(&scratch));
  const bool din = on_device(pcm_any), dout = on_device(out_any);
  const float* src = pcm_any;
  if (!din) {
    PTTS_CUDA_CHECK(cudaMemcpyAsync(scratch, pcm_any, (size_t)n * sizeof(float), cudaMemcpyHostToDevice, stream));
    src = reinterpret_cast<const float*>(scratch);
  }
  short* dst = dout ? out_any : reinterpret_cast<short*>(scratch + kInBytes);
  PTTS_CUDA_CHECK(launch_pcm_to_int16(src, dst, n, stream));
  if (launches) ++*launches;
  if (!dout) PTTS_CUDA_CHECK(cudaMemcpyAsync(out_any, dst, (size_t)n * sizeof(short), cudaMemcpyDeviceToHost, stream));
  PTTS_CUDA_CHECK(cudaStreamSynchronize(stream));
  return PTTS_OK;
}

}  // namespace ptts
