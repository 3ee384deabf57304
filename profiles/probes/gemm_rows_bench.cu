// Developer microbenchmark of the row-major tcgen05 GEMM (gemm_tc.cu) on the Mimi pass shapes at 64 streams x 4
// frames: the decoder transformer's four linears and the SEANet convolutions as overlapping-row GEMMs.  Includes
// the product kernel source directly, so it times exactly what the engine launches.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 --expt-relaxed-constexpr -o gemm_rows_bench.bin gemm_rows_bench.cu
#ifdef ROWS_TRACE_BUILD   // -DROWS_TRACE_BUILD: per-tile SM clock stamps of CTA 0 (producer / MMA issuer / epilogue), printed per shape
#define TC_ROWS_TRACE 1
#endif
#include "../../pocket_tts_b200/csrc/gemm_tc.cu"

#include <vector>

namespace ptts {
bool pdl_enabled() { return true; }
void prefer_max_smem_carveout(const void*) {}
}  // namespace ptts

using namespace ptts;

struct Shape { const char* name; int streams, rpb, valid, ld, K, N; int act; bool res; };

int main(int argc, char** argv) {
  const int B = 64, F = 4;
  std::vector<Shape> shapes = {
      {"mimi.qkv", 1, 1 << 30, 1 << 30, 512, 512, 1536, ACT_NONE, false},
      {"mimi.out_proj", 1, 1 << 30, 1 << 30, 512, 512, 512, ACT_NONE, true},
      {"mimi.ffn1", 1, 1 << 30, 1 << 30, 512, 512, 2048, ACT_GELU, false},
      {"ffn1 no act", 1, 1 << 30, 1 << 30, 512, 512, 2048, ACT_NONE, false},
      {"ffn1 N1536 gelu", 1, 1 << 30, 1 << 30, 512, 512, 1536, ACT_GELU, false},
      {"ffn1 N1024 gelu", 1, 1 << 30, 1 << 30, 512, 512, 1024, ACT_GELU, false},
      {"ffn1 N4096 none", 1, 1 << 30, 1 << 30, 512, 512, 4096, ACT_NONE, false},
      {"mimi.ffn2", 1, 1 << 30, 1 << 30, 2048, 2048, 512, ACT_NONE, true},
      {"seanet.conv0", B, 6 + 16 * F, 16 * F, 512, 3584, 512, ACT_ELU, false},
      {"seanet.convtr0", B, 1 + 16 * F, 16 * F, 512, 1024, 1536, ACT_ELU, false},
      {"seanet.res0.conv3", B, 2 + 96 * F, 96 * F, 256, 768, 128, ACT_ELU, false},
      {"seanet.res0.conv1", B, 96 * F, 96 * F, 128, 128, 256, ACT_ELU, true},
      {"seanet.convtr1", B, 1 + 96 * F, 96 * F, 256, 512, 640, ACT_ELU, false},
      {"seanet.res1.conv3", B, 2 + 480 * F, 480 * F, 128, 384, 64, ACT_ELU, false},
      {"seanet.res1.conv1", B, 480 * F, 480 * F, 64, 64, 128, ACT_ELU, true},
      {"seanet.convtr2", B, 1 + 480 * F, 480 * F, 128, 256, 256, ACT_ELU, false},
      {"seanet.res2.conv3", B, 2 + 1920 * F, 1920 * F, 64, 192, 64, ACT_ELU, false},
      {"seanet.res2.conv1", B, 1920 * F, 1920 * F, 64, 64, 64, ACT_ELU, true},
  };
  cudaStream_t st;
  cudaStreamCreate(&st);
  int n_sms = 148;
  cudaDeviceGetAttribute(&n_sms, cudaDevAttrMultiProcessorCount, 0);
  double total = 0;
  for (const Shape& s : shapes) {
    const bool plainA = s.rpb >= (1 << 30);
    const long long M = plainA ? (long long)B * 16 * F : (long long)s.streams * s.rpb;
    const long long a_elems = plainA ? M * s.ld : (long long)s.streams * s.rpb * s.ld + s.K;
    bf16 *A, *W, *out, *res;
    float* bias;
    cudaMalloc(&A, a_elems * 2 + 4096);
    cudaMalloc(&W, (size_t)s.N * s.K * 2);
    cudaMalloc(&out, (size_t)M * s.N * 2 + 4096);
    cudaMalloc(&res, (size_t)M * s.N * 2 + 4096);
    cudaMalloc(&bias, s.N * 4);
    cudaMemset(A, 0x3c, a_elems * 2);   // bf16 0x3c3c ~ 0.0115
    cudaMemset(W, 0x3c, (size_t)s.N * s.K * 2);
    cudaMemset(res, 0, (size_t)M * s.N * 2);
    cudaMemset(bias, 0, s.N * 4);
    GemmProblem p;
    p.A = A; p.W = W; p.M = (int)M; p.N = s.N; p.K = s.K;
    p.a_map = plainA ? RowMap{0, 1 << 30, 1 << 30, s.ld} : RowMap{(long long)s.rpb * s.ld, s.rpb, s.valid, s.ld};
    p.ep.bias = bias; p.ep.act_post = s.act; p.ep.out = out;
    p.ep.out_map = plainA ? RowMap{0, 1 << 30, 1 << 30, s.N} : RowMap{(long long)s.valid * s.N, s.rpb, s.valid, s.N};
    if (s.res) { p.ep.res = res; p.ep.res_map = p.ep.out_map; }
    const long long a_rows = plainA ? M : M - (s.K / s.ld - 1);
    const int iters = 20;
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    cudaError_t err = cudaSuccess;
    for (int i = 0; i < 3 && err == cudaSuccess; ++i) err = gemm_tc_launch(p, 1, a_rows, n_sms, false, st);
    cudaStreamSynchronize(st);
    cudaEventRecord(e0, st);
    for (int i = 0; i < iters && err == cudaSuccess; ++i) err = gemm_tc_launch(p, 1, a_rows, n_sms, false, st);
    cudaEventRecord(e1, st);
    cudaError_t e2 = cudaStreamSynchronize(st);
    if (err != cudaSuccess || e2 != cudaSuccess) { printf("%s failed: %s %s\n", s.name, cudaGetErrorString(err), cudaGetErrorString(e2)); return 1; }
    float ms = 0;
    cudaEventElapsedTime(&ms, e0, e1);
    const double us = ms * 1e3 / iters;
    const long long valid_rows = plainA ? M : (long long)s.streams * s.valid;
    const double tf = 2.0 * valid_rows * s.N * s.K / (us * 1e-6) / 1e12;
    printf("%-20s M %7lld N %5d K %5d: %7.1f us  %7.1f TFLOP/s\n", s.name, M, s.N, s.K, us, tf);
#ifdef TC_ROWS_TRACE
    {
      static long long h[16 + 8 * 32];
      cudaMemcpyFromSymbol(h, g_rows_trace, sizeof h);
      const long long t0 = h[0];
      for (int t = 0; t < 32 && h[16 + t * 8 + 5]; ++t) {
        const long long* r = h + 16 + t * 8;
        printf("   tile %d (clk since start): load first/last issue %6lld %6lld | mma acc-free %6lld first-operands %6lld last-kb %6lld | epi acc-ready %6lld done %6lld\n",
               t, r[0] - t0, r[1] - t0, r[2] - t0, r[3] - t0, r[4] - t0, r[5] - t0, r[6] - t0);
      }
      static long long e[16 * 32];
      int ntile = 0;
      cudaMemcpyFromSymbol(e, g_epi_trace, sizeof e);
      cudaMemcpyFromSymbol(&ntile, g_epi_tile, sizeof ntile);
      for (int t = 0; t < 2 && t < ntile; ++t) {
        const long long* r = e + t * 16;
        printf("   epilogue of launch-tile %d, warp 2 (clk): entry 0 | tables+barrier %lld | chunk A: tmem_ld %lld staged %lld rows done %lld | chunk B: tmem_ld %lld staged %lld rows done %lld\n",
               t, r[1] - r[0], r[2] - r[0], r[3] - r[0], r[4] - r[0], r[5] - r[0], r[6] - r[0], r[7] - r[0]);
      }
      static long long z[16 + 8 * 32];
      static long long ze[16 * 32];
      const int zero = 0;
      cudaMemcpyToSymbol(g_rows_trace, z, sizeof z);
      cudaMemcpyToSymbol(g_epi_trace, ze, sizeof ze);
      cudaMemcpyToSymbol(g_epi_tile, &zero, sizeof zero);
    }
#endif
    total += us;
    cudaFree(A); cudaFree(W); cudaFree(out); cudaFree(res); cudaFree(bias);
  }
  printf("sum %.1f us per 4-frame pass = %.1f us/frame (transformer linears counted once; x2 layers in the engine)\n", total, total / F);
  return 0;
}
