// Developer microbenchmark + phase trace of the fused last SEANet stage (seanet_fused.cu) on synthetic data:
// 64 streams x {1, 2, 4} frames, one tile per CTA and persistent forms.  Includes the product sources directly.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 --expt-relaxed-constexpr -o seanet_fused_bench.bin seanet_fused_bench.cu -lcuda
#include "../../pocket_tts_b200/csrc/gemm_tc.cu"
#include "../../pocket_tts_b200/csrc/seanet_fused.cu"

#include <vector>

namespace ptts {
bool pdl_enabled() { return true; }
void prefer_max_smem_carveout(const void*) {}
}  // namespace ptts
using namespace ptts;

int main() {
  const int n = 64;
  cudaStream_t st;
  cudaStreamCreate(&st);
  bf16 *wtr, *w3, *w1;
  cudaMalloc(&wtr, 256 * 256 * 2); cudaMalloc(&w3, 64 * 192 * 2); cudaMalloc(&w1, 64 * 64 * 2);
  cudaMemset(wtr, 0x3c, 256 * 256 * 2); cudaMemset(w3, 0x3c, 64 * 192 * 2); cudaMemset(w1, 0x3c, 64 * 64 * 2);
  FusedStage2Consts cst{};
  for (int j = 0; j < 192; ++j) cst.w_last[j] = 0.01f;
  unsigned long long* trace;
  cudaMalloc(&trace, 16 * 12 * 8);
  for (int frames : {1, 2, 4}) {
    const int T = 480 * frames;
    bf16 *x, *res, *last;
    float* pcm;
    cudaMalloc(&x, (size_t)n * (1 + T) * 128 * 2 + 4096);
    cudaMalloc(&res, (size_t)n * (2 + 4 * T) * 64 * 2);
    cudaMalloc(&last, (size_t)n * (2 + 4 * T) * 64 * 2);
    cudaMalloc(&pcm, (size_t)n * 4 * T * 4);
    cudaMemset(x, 0x3c, (size_t)n * (1 + T) * 128 * 2);
    cudaMemset(res, 0, (size_t)n * (2 + 4 * T) * 64 * 2);
    cudaMemset(last, 0, (size_t)n * (2 + 4 * T) * 64 * 2);
    for (int cap : {1 << 30, 148}) {
      cudaEvent_t e0, e1;
      cudaEventCreate(&e0); cudaEventCreate(&e1);
      cudaError_t err = cudaSuccess;
      for (int i = 0; i < 3; ++i) err = launch_seanet_stage2_fused(x, wtr, w3, w1, cst, res, last, pcm, n, T, cap, false, st);
      cudaStreamSynchronize(st);
      cudaEventRecord(e0, st);
      const int iters = 20;
      for (int i = 0; i < iters; ++i) err = launch_seanet_stage2_fused(x, wtr, w3, w1, cst, res, last, pcm, n, T, cap, false, st);
      cudaEventRecord(e1, st);
      cudaError_t e2 = cudaStreamSynchronize(st);
      if (err != cudaSuccess || e2 != cudaSuccess) { printf("failed: %s %s\n", cudaGetErrorString(err), cudaGetErrorString(e2)); return 1; }
      float ms;
      cudaEventElapsedTime(&ms, e0, e1);
      const int tiles = n * ((T + 126) / 127);
      printf("frames %d, %s: %.1f us per pass (%d tiles, %.2f us per tile-slot)\n", frames, cap == 148 ? "persistent 148" : "tile per CTA", ms * 1e3 / iters,
             tiles, ms * 1e3 / iters / ((tiles + 147) / 148));
      if (cap == 148 && frames == 4) {
        cudaMemset(trace, 0, 16 * 12 * 8);
        launch_seanet_stage2_fused(x, wtr, w3, w1, cst, res, last, pcm, n, T, cap, false, st, trace);
        cudaStreamSynchronize(st);
        unsigned long long h[16 * 12];
        cudaMemcpy(h, trace, sizeof h, cudaMemcpyDeviceToHost);
        const char* names[12] = {"mma:tmem_free", "mma:A issued", "epi:da_full", "epi:P0-3 written", "epi:b_full3", "epi:shifted written",
                                 "epi:epiB done", "epi:tmem drained", "epi:tile end", "mma:B2,B3 issued", "mma:all issued", ""};
        for (int t = 0; t < 7; ++t) {
          printf("tile %d:", t);
          for (int s = 0; s < 11; ++s) printf("  %s %+.2f", names[s], h[t * 12 + s] ? (double)((long long)h[t * 12 + s] - (long long)h[t * 12]) / 1e3 : -1.0);
          printf("\n");
        }
      }
    }
    cudaFree(x); cudaFree(res); cudaFree(last); cudaFree(pcm);
  }
  return 0;
}
