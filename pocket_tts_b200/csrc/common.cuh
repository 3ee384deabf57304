// Shared device helpers for the pocket_tts_b200 kernels (sm_100a only).
#pragma once
#include <cuda_bf16.h>
#include <cuda_fp8.h>
#include <cuda_runtime.h>
#include <stdint.h>

namespace ptts {

typedef __nv_bfloat16 bf16;
// FP8 (e4m3) KV-cache element (PTTS kv_fp8 mode): one byte, paired with an fp32 scale per cached row
struct fp8e4 { uint8_t b; };

// ---- activation codes used by GEMM epilogues
enum Act : int { ACT_NONE = 0, ACT_GELU = 1, ACT_SILU = 2, ACT_ELU = 3 };

__device__ __forceinline__ float gelu_erf(float x) {
  // F.gelu default (erf form), mimi_transformer.py:42
  return 0.5f * x * (1.0f + erff(x * 0.70710678118654752440f));
}
__device__ __forceinline__ float silu_f(float x) { return x / (1.0f + expf(-x)); }
__device__ __forceinline__ float elu_f(float x) { return x > 0.0f ? x : expm1f(x); }  // alpha = 1

__device__ __forceinline__ float apply_act(float v, int act) {
  switch (act) {
    case ACT_GELU: return gelu_erf(v);
    case ACT_SILU: return silu_f(v);
    case ACT_ELU: return elu_f(v);
    default: return v;
  }
}

// ---- storage-type helpers (float or bf16 activations / weights)
__device__ __forceinline__ float to_f(float v) { return v; }
__device__ __forceinline__ float to_f(bf16 v) { return __bfloat162float(v); }
template <typename T> __device__ __forceinline__ T from_f(float v);
template <> __device__ __forceinline__ float from_f<float>(float v) { return v; }
template <> __device__ __forceinline__ bf16 from_f<bf16>(float v) { return __float2bfloat16_rn(v); }

// load 4 consecutive elements as float4 (pointer must be aligned to 4 elements)
__device__ __forceinline__ float4 load4(const float* p) { return *reinterpret_cast<const float4*>(p); }
__device__ __forceinline__ float4 load4(const bf16* p) {
  uint2 r = *reinterpret_cast<const uint2*>(p);
  __nv_bfloat162 a = *reinterpret_cast<__nv_bfloat162*>(&r.x);
  __nv_bfloat162 b = *reinterpret_cast<__nv_bfloat162*>(&r.y);
  float2 fa = __bfloat1622float2(a), fb = __bfloat1622float2(b);
  return make_float4(fa.x, fa.y, fb.x, fb.y);
}
__device__ __forceinline__ void store4(float* p, float4 v) { *reinterpret_cast<float4*>(p) = v; }
__device__ __forceinline__ void store4(bf16* p, float4 v) {
  __nv_bfloat162 a = __floats2bfloat162_rn(v.x, v.y);
  __nv_bfloat162 b = __floats2bfloat162_rn(v.z, v.w);
  uint2 r;
  r.x = *reinterpret_cast<uint32_t*>(&a);
  r.y = *reinterpret_cast<uint32_t*>(&b);
  *reinterpret_cast<uint2*>(p) = r;
}
__device__ __forceinline__ float2 load2(const float* p) { return *reinterpret_cast<const float2*>(p); }
__device__ __forceinline__ float2 load2(const bf16* p) {
  return __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(p));
}
__device__ __forceinline__ float2 load2(const fp8e4* p) {   // two e4m3 values (unscaled)
  const __half2_raw h = __nv_cvt_fp8x2_to_halfraw2(*reinterpret_cast<const __nv_fp8x2_storage_t*>(p), __NV_E4M3);
  return __half22float2(*reinterpret_cast<const __half2*>(&h));
}
__device__ __forceinline__ void store2(fp8e4* p, float2 v) {
  *reinterpret_cast<__nv_fp8x2_storage_t*>(p) = __nv_cvt_float2_to_fp8x2(v, __NV_SATFINITE, __NV_E4M3);
}
__device__ __forceinline__ void store2(float* p, float2 v) { *reinterpret_cast<float2*>(p) = v; }
__device__ __forceinline__ void store2(bf16* p, float2 v) {
  *reinterpret_cast<__nv_bfloat162*>(p) = __floats2bfloat162_rn(v.x, v.y);
}

// ---- programmatic dependent launch: `pdl_wait` blocks until the previous kernel of the stream has
// completed and its writes are visible (a no-op when launched without the attribute); every kernel
// of the frame step calls it before its first dependent global access.
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void pdl_launch_dependents() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }

// ---- developer timeline (PTTS_TIMELINE=1): thread 0 of CTA (0,0,0) of every instrumented kernel records its
// entry, the return of its programmatic-dependent-launch wait and its exit (globaltimer, ns) into a ring of
// records; the engine prints the last records when it is destroyed.  Each translation unit has its own copy
// of the buffer pointer (set through tl_enable_<unit>), so no relocatable device code is needed.
constexpr unsigned TL_MAX = 4096;        // records; buffer = 2 header words + TL_W words per record
constexpr unsigned TL_W = 8;             // kind, entry, wait returned, exit, 4 kernel-specific marks
static __device__ unsigned long long* tl_buf_dev = nullptr;
__device__ __forceinline__ unsigned long long tl_now() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  return t;
}
__device__ __forceinline__ unsigned tl_begin(int kind) {
  unsigned long long* b = tl_buf_dev;
  if (!b || threadIdx.x != 0 || (blockIdx.x | blockIdx.y | blockIdx.z) != 0) return ~0u;
  const unsigned s = (unsigned)(atomicAdd(b, 1ull) % TL_MAX);
  b[2 + TL_W * s] = (unsigned long long)kind;
  b[2 + TL_W * s + 1] = tl_now();
#pragma unroll
  for (unsigned i = 2; i < TL_W; ++i) b[2 + TL_W * s + i] = 0;
  return s;
}
__device__ __forceinline__ void tl_mark(unsigned s, int which) {   // which: 2 = wait returned, 3 = exit, 4..7 free
  if (s != ~0u) tl_buf_dev[2 + TL_W * s + which] = tl_now();
}
#define PTTS_TL_ENABLE_FN(unit) \
  void tl_enable_##unit(unsigned long long* p) { cudaMemcpyToSymbol(tl_buf_dev, &p, sizeof p); }

// ---- reductions
__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}
// block-wide sum for blockDim.x <= 1024 (multiple of 32); `scratch` holds >= 33 floats
__device__ __forceinline__ float block_sum(float v, float* scratch) {
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5, nw = (blockDim.x + 31) >> 5;
  v = warp_sum(v);
  __syncthreads();  // protect scratch reuse
  if (lane == 0) scratch[w] = v;
  __syncthreads();
  float r = (lane < nw) ? scratch[lane] : 0.0f;
  r = warp_sum(r);
  return r;
}

// ---- generic "batched rows" addressing used by GEMM operands / outputs:
// logical row m -> (i, t) = divmod(m, rows_per_batch); valid iff t < valid_rows;
// address = base + i*batch_stride + t*ld (+ column).
struct RowMap {
  long long batch_stride;  // elements
  int rows_per_batch;      // >= 1
  int valid_rows;          // rows [0, valid_rows) of each batch are real outputs
  int ld;                  // elements between consecutive rows
};

// ---- host-side launch helper: every kernel of the frame step is launched with the programmatic
// stream serialization attribute (PTTS_PDL=0 disables) so that its prologue overlaps the tail of
// the previous kernel; kernels call pdl_wait() before their first dependent access.
bool pdl_enabled();

// Per-process caches of per-device settings (cudaFuncSetAttribute opt-ins): first() is true once per CUDA device.
struct DeviceOnce {
  unsigned long long mask = 0;
  bool first() {
    int d = 0;
    cudaGetDevice(&d);
    const unsigned long long bit = 1ull << (d & 63);
    if (mask & bit) return false;
    mask |= bit;
    return true;
  }
};
// Every kernel asks for the same (maximum) shared-memory carveout: alternating carveouts between
// consecutive kernels makes the SMs drain and repartition L1/shared memory at every boundary.
void prefer_max_smem_carveout(const void* kern);
template <typename... KArgs, typename... Args>
inline cudaError_t launch_k(void (*kern)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t st, Args&&... args) {
  prefer_max_smem_carveout(reinterpret_cast<const void*>(kern));
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = grid;
  cfg.blockDim = block;
  cfg.dynamicSmemBytes = smem;
  cfg.stream = st;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = pdl_enabled() ? 1 : 0;
  return cudaLaunchKernelEx(&cfg, kern, static_cast<KArgs>(args)...);
}

}  // namespace ptts

// ---- host-side error plumbing
#define PTTS_CUDA_CHECK(expr)                                                        \
  do {                                                                               \
    cudaError_t _e = (expr);                                                         \
    if (_e != cudaSuccess) {                                                         \
      ptts::set_error_cuda(_e, #expr, __FILE__, __LINE__);                           \
      return PTTS_ECUDA;                                                             \
    }                                                                                \
  } while (0)
