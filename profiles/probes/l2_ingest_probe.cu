// Probe: how fast can ONE SM pull L2-resident data into shared memory with bulk async copies (the TMA data path)?
// The stream-split flow-head kernel streams all 6.3 MB of head weights through every CTA, so this number bounds it.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o l2_ingest_probe l2_ingest_probe.cu && ./l2_ingest_probe
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdint>
#include <vector>

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

template <int STAGES, int CHUNK, int CL>
__global__ void __launch_bounds__(64, 1) ingest_kernel(const char* src, size_t bytes, int reps, unsigned long long* out_ns) {
  extern __shared__ __align__(1024) char sm[];
  __shared__ uint64_t full[STAGES];
  const int tid = threadIdx.x;
  uint32_t rank = 0;
  if (CL > 1) asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(rank));
  if (tid == 0) {
    for (int s = 0; s < STAGES; ++s)
      asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(&full[s])), "r"(1));
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();
  if (CL > 1) {
    asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
    asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
  }
  const size_t n_chunks = bytes / CHUNK;
  unsigned long long t0 = 0, t1 = 0;
  if (tid == 0) {
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t0));
    size_t issued = 0, done = 0;
    const size_t total = n_chunks * reps;
    uint32_t ph[STAGES] = {};
    // prime
    auto issue = [&](size_t i) {
      const int s = i % STAGES;
      const size_t c = i % n_chunks;
      const uint32_t bar = smem_u32(&full[s]);
      asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(CHUNK) : "memory");
      if (CL == 1) {
        asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                     ::"r"(smem_u32(sm + (size_t)s * CHUNK)), "l"(src + c * CHUNK), "r"(CHUNK), "r"(bar) : "memory");
      } else {
        // each CTA of the cluster fetches 1/CL of the chunk and multicasts it to every CTA
        const uint32_t part = CHUNK / CL;
        asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes.multicast::cluster [%0], [%1], %2, [%3], %4;"
                     ::"r"(smem_u32(sm + (size_t)s * CHUNK + rank * part)), "l"(src + c * CHUNK + rank * part), "r"(part), "r"(bar),
                       "h"((uint16_t)((1u << CL) - 1)) : "memory");
      }
    };
    for (; issued < STAGES && issued < total; ++issued) issue(issued);
    while (done < total) {
      const int s = done % STAGES;
      uint32_t ok = 0;
      for (unsigned spin = 0; !ok && spin < (1u << 24); ++spin)
        asm volatile("{.reg .pred p; mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2; selp.u32 %0,1,0,p;}" : "=r"(ok) : "r"(smem_u32(&full[s])), "r"(ph[s]) : "memory");
      if (!ok) { out_ns[blockIdx.x] = ~0ull; break; }   // protocol slip of the unsynchronised multicast variant: give up
      ph[s] ^= 1;
      ++done;
      if (CL > 1) {   // peers must have consumed the stage too before anyone overwrites it: keep it simple, cluster barrier by thread 0 is not possible; see below
      }
      if (issued < total) { issue(issued); ++issued; }
    }
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t1));
    if (done == total) out_ns[blockIdx.x] = t1 - t0;
  }
  __syncthreads();
  if (CL > 1) {
    asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
    asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
  }
}

template <int STAGES, int CHUNK, int CL>
void run(const char* d, size_t bytes, int ctas, int reps, unsigned long long* d_ns) {
  auto k = ingest_kernel<STAGES, CHUNK, CL>;
  const size_t smem = (size_t)STAGES * CHUNK;
  cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = dim3(ctas); cfg.blockDim = dim3(64); cfg.dynamicSmemBytes = smem;
  cudaLaunchAttribute at[1];
  at[0].id = cudaLaunchAttributeClusterDimension; at[0].val.clusterDim.x = CL; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
  cfg.attrs = at; cfg.numAttrs = 1;
  for (int it = 0; it < 3; ++it) {
    cudaError_t e = cudaLaunchKernelEx(&cfg, k, d, bytes, reps, d_ns);
    if (e != cudaSuccess) { printf("launch failed: %s\n", cudaGetErrorString(e)); return; }
    cudaDeviceSynchronize();
  }
  std::vector<unsigned long long> ns(ctas);
  cudaMemcpy(ns.data(), d_ns, ctas * sizeof(unsigned long long), cudaMemcpyDeviceToHost);
  unsigned long long mx = 0; double avg = 0;
  for (auto v : ns) { mx = v > mx ? v : mx; avg += v; }
  avg /= ctas;
  const double gb = (double)bytes * reps / 1e9;
  printf("stages %2d chunk %5d cluster %d ctas %3d: per-CTA %.1f GB/s (slowest CTA %.1f us for %.0f KB; avg %.1f GB/s)  %.1f B/clk @1.965GHz\n", STAGES, CHUNK, CL,
         ctas, gb / (mx * 1e-9), mx / 1e3, bytes * reps / 1024.0, gb / (avg * 1e-9), gb / (mx * 1e-9) / 1.965);
}

int main() {
  const size_t bytes = 512 * 1024;   // one 512 x 512 bf16 matrix
  char* d; unsigned long long* d_ns;
  cudaMalloc(&d, 16 * bytes); cudaMemset(d, 1, 16 * bytes);
  cudaMalloc(&d_ns, 1024 * 8);
  const int reps = 12;
  for (int ctas : {1, 4, 8, 16, 32, 148}) {
    run<8, 16384, 1>(d, bytes, ctas, reps, d_ns);
    run<12, 16384, 1>(d, bytes, ctas, reps, d_ns);
    run<6, 32768, 1>(d, bytes, ctas, reps, d_ns);
    run<24, 8192, 1>(d, bytes, ctas, reps, d_ns);
  }
  // NOTE: multicast variants reuse stages without a consumer handshake between CTAs: a timing probe only
  for (int ctas : {4, 8, 16, 32}) {
    run<12, 16384, 2>(d, bytes, ctas, reps, d_ns);
    run<12, 16384, 4>(d, bytes, ctas, reps, d_ns);
  }
  return 0;
}
