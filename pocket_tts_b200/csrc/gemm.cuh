// GEMM front-end shared by the SIMT (fp32-exact) and tcgen05 (bf16) back-ends.
//
//   D[m, n] = epilogue( sum_k A[m, k] * W[n, k] )        A, W both K-major ("TN")
//
// A is addressed through a RowMap so that one kernel covers plain linears, streaming
// Conv1d (row t of a stream's channels-last buffer is the K*C_in contiguous elements that
// start at time t: rows overlap, ld = C_in) and K=2S ConvTranspose1d (rows of 2*C_in,
// ld = C_in, N = S*C_out) — see DESIGN.md "convolutions as overlapping-row GEMMs".
#pragma once
#include <cuda.h>

#include "common.cuh"

namespace ptts {

struct GemmEpilogue {
  const float* bias = nullptr;      // [N]
  const float* colscale = nullptr;  // [N]   LayerScale (layer_scale.py:10-11)
  const float* gate = nullptr;      // [M, gate_ld] AdaLN gate (mlp.py:107-111)
  int gate_ld = 0;
  const void* res = nullptr;        // residual, same element type as `out`
  RowMap res_map{0, 1 << 30, 1 << 30, 0};
  float alpha = 1.0f;               // v = alpha * gate * colscale * act_pre(acc + bias) + res
  int act_pre = ACT_NONE;
  int act_post = ACT_NONE;          // out = act_post(v); out2 = v
  void* out = nullptr;
  RowMap out_map{0, 1 << 30, 1 << 30, 0};
  void* out2 = nullptr;             // optional pre-act_post copy
  RowMap out2_map{0, 1 << 30, 1 << 30, 0};
  // LayerNorm folded into a producer / consumer GEMM pair (skinny tcgen05 path only):
  //  producer (residual GEMM writing the fp32 stream h): also stores v as bf16 in ln_raw [M, N] and, per row and
  //  per block of 16 features, (sum, M2) of v in ln_part_out [M][N/16];
  //  consumer (GEMM on ln_raw with gamma folded into the weights): acc -> rstd * (acc - mean * ln_c1[n]) + ln_c2[n]
  //  before bias / activation, with mean / rstd of each row combined from ln_part_in [M][ln_nparts].
  bf16* ln_raw = nullptr;
  float2* ln_part_out = nullptr;
  const float2* ln_part_in = nullptr;
  int ln_nparts = 0;                // partials per row (consumer K / 16)
  float ln_eps = 1e-5f;
  const float* ln_c1 = nullptr;     // [N] sum_k bf16(W[n,k] gamma[k])
  const float* ln_c2 = nullptr;     // [N] sum_k W[n,k] beta[k]
};

// Optional side job of a skinny GEMM: while it is resident ahead of its producer (programmatic dependent launch),
// its epilogue warps pull the cached K/V rows that the NEXT kernel (decode attention of the same layer) will read
// into L2.  Region r of stream i starts at base + slot_ids[i] * slot_stride + r * region_stride and holds
// offsets[slot] rows of row_bytes.
struct KvPrefetchJob {
  const char* base = nullptr;
  const int* slot_ids = nullptr;
  const int* offsets = nullptr;
  long long slot_stride = 0, region_stride = 0;
  int row_bytes = 0, regions = 0, n = 0;
};

struct GemmProblem {
  const void* A = nullptr;
  RowMap a_map{0, 1 << 30, 1 << 30, 0};
  const void* W = nullptr;  // [N, K] row-major
  int M = 0, N = 0, K = 0;
  GemmEpilogue ep;
  int max_splits = 8;       // skinny tcgen05 path: upper bound of the cluster split-K factor
  KvPrefetchJob kv_prefetch;
};

// element types: 0 = float, 1 = bf16
cudaError_t gemm_simt_launch(const GemmProblem& p, int in_type, int out_type, cudaStream_t st);
// bf16 x bf16 -> fp32 on tcgen05 tensor cores (operands via TMA); a_rows_alloc = addressable A rows
void tc_trace_dump(const char* what);
// 2-D bf16 SWIZZLE_128B tensor map (inner dim K, box 64 x box_rows); cached per argument tuple
bool tc_make_map(CUtensorMap* out, const void* base, long long rows, int K, long long ld, int box_rows);
cudaError_t gemm_tc_launch(const GemmProblem& p, int out_type, long long a_rows_alloc, int n_sms, bool pdl,
                           cudaStream_t st);

// shared epilogue math (used by both back-ends): returns v (pre act_post)
template <typename TO>
__device__ __forceinline__ void epilogue_store4(const GemmEpilogue& ep, int m, int n0, int N,
                                                float acc[4]) {
  // row validity / addresses
  const int ob = m / ep.out_map.rows_per_batch, ot = m - ob * ep.out_map.rows_per_batch;
  if (ot >= ep.out_map.valid_rows) return;
  float v[4];
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    const int n = n0 + j;
    float x = acc[j];
    if (n < N) {
      if (ep.bias) x += ep.bias[n];
      x = apply_act(x, ep.act_pre);
      if (ep.gate) x *= ep.gate[(long long)m * ep.gate_ld + n];
      if (ep.colscale) x *= ep.colscale[n];
      x *= ep.alpha;
    }
    v[j] = x;
  }
  if (ep.res) {
    const int rb = m / ep.res_map.rows_per_batch, rt = m - rb * ep.res_map.rows_per_batch;
    const TO* r = reinterpret_cast<const TO*>(ep.res) + rb * ep.res_map.batch_stride +
                  (long long)rt * ep.res_map.ld + n0;
    if (n0 + 3 < N) {
      float4 rv = load4(r);
      v[0] += rv.x; v[1] += rv.y; v[2] += rv.z; v[3] += rv.w;
    } else {
      for (int j = 0; j < 4 && n0 + j < N; ++j) v[j] += to_f(r[j]);
    }
  }
  if (ep.out2) {
    const int b2 = m / ep.out2_map.rows_per_batch, t2 = m - b2 * ep.out2_map.rows_per_batch;
    TO* o2 = reinterpret_cast<TO*>(ep.out2) + b2 * ep.out2_map.batch_stride +
             (long long)t2 * ep.out2_map.ld + n0;
    if (n0 + 3 < N) store4(o2, make_float4(v[0], v[1], v[2], v[3]));
    else for (int j = 0; j < 4 && n0 + j < N; ++j) o2[j] = from_f<TO>(v[j]);
  }
  TO* o = reinterpret_cast<TO*>(ep.out) + ob * ep.out_map.batch_stride +
          (long long)ot * ep.out_map.ld + n0;
  if (n0 + 3 < N) {
    store4(o, make_float4(apply_act(v[0], ep.act_post), apply_act(v[1], ep.act_post),
                          apply_act(v[2], ep.act_post), apply_act(v[3], ep.act_post)));
  } else {
    for (int j = 0; j < 4 && n0 + j < N; ++j) o[j] = from_f<TO>(apply_act(v[j], ep.act_post));
  }
}

}  // namespace ptts
