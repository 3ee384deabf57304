// Probe 2: per-SM L2 -> shared-memory throughput of bulk async copies, 1-D vs 2-D tensor (SWIZZLE_128B) form, one
// or several issuing warps.  (Probe 1 showed ~0.3 us per 1-D bulk op regardless of its size with one issuing thread.)
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o l2_ingest_probe2.bin l2_ingest_probe2.cu -lcuda
#include <cuda.h>
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdint>
#include <vector>

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

// NW issuing warps; warp w owns stages s with s % NW == w.  TENSOR: 2-D tiled loads of box 64 x ROWS bf16.
template <int STAGES, int ROWS, int NW, bool TENSOR, bool LANES = false>
__global__ void __launch_bounds__(NW * 32, 1) ingest2(const char* src, const __grid_constant__ CUtensorMap map, int n_tiles, int reps,
                                                      unsigned long long* out_ns) {
  extern __shared__ __align__(1024) char smraw[];
  char* sm = (char*)(((uintptr_t)smraw + 1023) & ~(uintptr_t)1023);
  __shared__ uint64_t full[STAGES];
  constexpr uint32_t CHUNK = ROWS * 128;
  const int tid = threadIdx.x;
  const int warp = LANES ? (tid < NW ? tid : 0) : (tid >> 5);
  const int lane = LANES ? (tid < NW ? 0 : 1) : (tid & 31);
  if (tid == 0)
    for (int s = 0; s < STAGES; ++s) asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(&full[s])), "r"(1));
  asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  __syncthreads();
  unsigned long long t0 = 0, t1 = 0;
  if (lane == 0) {
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t0));
    const int total = n_tiles * reps;       // ops overall; this warp takes i % NW == warp
    auto issue = [&](int i) {
      const int s = i % STAGES, c = i % n_tiles;
      const uint32_t bar = smem_u32(&full[s]);
      asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(CHUNK) : "memory");
      if (TENSOR) {
        // tile c: rows (c / 8) * ROWS .., k-block c % 8 of a [rows, 512] bf16 matrix
        asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
                     ::"r"(smem_u32(sm + (size_t)s * CHUNK)), "l"(&map), "r"(bar), "r"((c % 8) * 64), "r"((c / 8) * ROWS) : "memory");
      } else {
        asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                     ::"r"(smem_u32(sm + (size_t)s * CHUNK)), "l"(src + (size_t)c * CHUNK), "r"(CHUNK), "r"(bar) : "memory");
      }
    };
    int issued = warp, done = warp;
    for (; issued < STAGES && issued < total; issued += NW) issue(issued);
    bool bad = false;
    while (done < total) {
      const int s = done % STAGES;
      const uint32_t par = (done / STAGES) & 1;
      uint32_t ok = 0;
      for (unsigned spin = 0; !ok && spin < (1u << 24); ++spin)
        asm volatile("{.reg .pred p; mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2; selp.u32 %0,1,0,p;}" : "=r"(ok) : "r"(smem_u32(&full[s])), "r"(par) : "memory");
      if (!ok) { bad = true; break; }
      done += NW;
      if (issued < total) { issue(issued); issued += NW; }
    }
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t1));
    if (warp == 0) out_ns[blockIdx.x] = bad ? ~0ull : t1 - t0;
  }
  __syncthreads();
}

typedef CUresult (*EncodeFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*, const cuuint32_t*,
                             const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

template <int STAGES, int ROWS, int NW, bool TENSOR, bool LANES = false>
void run(const char* d, EncodeFn enc, int ctas, int reps, unsigned long long* d_ns) {
  const size_t bytes = 512 * 1024;
  const int n_tiles = (int)(bytes / (ROWS * 128));
  CUtensorMap map{};
  cuuint64_t dims[2] = {512, 512}, strides[1] = {1024};
  cuuint32_t box[2] = {64, (cuuint32_t)ROWS}, es[2] = {1, 1};
  CUresult r = enc(&map, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, (void*)d, dims, strides, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
                   CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) { printf("encode failed %d\n", (int)r); return; }
  auto k = ingest2<STAGES, ROWS, NW, TENSOR, LANES>;
  const size_t smem = (size_t)STAGES * ROWS * 128 + 1024;
  cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  for (int it = 0; it < 3; ++it) {
    k<<<ctas, NW * 32, smem>>>(d, map, n_tiles, reps, d_ns);
    cudaError_t e = cudaDeviceSynchronize();
    if (e != cudaSuccess) { printf("kernel failed: %s\n", cudaGetErrorString(e)); return; }
  }
  std::vector<unsigned long long> ns(ctas);
  cudaMemcpy(ns.data(), d_ns, ctas * sizeof(unsigned long long), cudaMemcpyDeviceToHost);
  unsigned long long mx = 0;
  for (auto v : ns) mx = v > mx ? v : mx;
  const double gb = (double)bytes * reps / 1e9;
  printf("%s%s stages %2d box 64x%3d (%2d KB) warps %d ctas %3d: %.1f GB/s per CTA (%.1f us for %d x 512 KB), %.1f B/clk\n", TENSOR ? "tensor2d" : "bulk1d  ", LANES ? " lanes-of-one-warp" : "",
         STAGES, ROWS, ROWS / 8, NW, ctas, gb / (mx * 1e-9), mx / 1e3, reps, gb / (mx * 1e-9) / 1.965);
}

int main() {
  EncodeFn enc = nullptr;
  cudaDriverEntryPointQueryResult q;
  cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", (void**)&enc, cudaEnableDefault, &q);
  if (!enc) { printf("no cuTensorMapEncodeTiled\n"); return 1; }
  char* d; unsigned long long* d_ns;
  cudaMalloc(&d, 4 << 20); cudaMemset(d, 1, 4 << 20);
  cudaMalloc(&d_ns, 1024 * 8);
  const int reps = 12;
  for (int ctas : {1, 148}) {
    run<12, 128, 1, false>(d, enc, ctas, reps, d_ns);
    run<12, 128, 2, false>(d, enc, ctas, reps, d_ns);
    run<12, 128, 4, false>(d, enc, ctas, reps, d_ns);
    run<12, 128, 1, true>(d, enc, ctas, reps, d_ns);
    run<12, 128, 2, true>(d, enc, ctas, reps, d_ns);
    run<12, 128, 4, true>(d, enc, ctas, reps, d_ns);
    run<6, 256, 1, true>(d, enc, ctas, reps, d_ns);
    run<6, 256, 2, true>(d, enc, ctas, reps, d_ns);
    run<24, 64, 4, true>(d, enc, ctas, reps, d_ns);
    run<24, 64, 1, true>(d, enc, ctas, reps, d_ns);
    run<12, 128, 2, true, true>(d, enc, ctas, reps, d_ns);
    run<12, 128, 4, true, true>(d, enc, ctas, reps, d_ns);
    run<12, 128, 8, true>(d, enc, ctas, reps, d_ns);
    run<6, 256, 4, true>(d, enc, ctas, reps, d_ns);
  }
  return 0;
}
