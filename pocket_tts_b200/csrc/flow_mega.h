// Persistent FlowLM frame-step kernel: the whole FlowLM decode step (flow_lm.py:96-157,
// mimi_transformer.py:12-54, transformer.py:135-158) and the flow head / LSD sampler
// (flow_lm.py:19-40, mlp.py:134-215) of a batch of <= 128 streams as ONE launch.
//
// The step is a program of phases separated by grid barriers.  Three phase kinds:
//   GEMM  items = (feature tile, K split) of D[stream, feature] = A W^T on tcgen05 (A: the bf16
//         activation rows via TMA, W: N_tile <= 64 weight rows via TMA, fp32 accumulator in TMEM).
//         Split-K items store raw fp32 partial slabs that the next (row-wise / attention) phase sums
//         in a fixed order; full-K items apply bias + activation and store the next GEMM's operand.
//   ROWS  one warp per stream row: residual add + LayerNorm, out_norm + EOS, AdaLN modulate, Euler
//         update ... (everything that needs a whole row).
//   ATTN  RoPE + KV append + one-query attention per (stream, head).
// The TMA producer of every CTA runs ahead of the barriers with its WEIGHT tiles (they depend on
// nothing), so the weight stream of phase p+1 is in flight while phase p computes; only the
// activation tiles wait for the grid barrier.
#pragma once
#include <cuda.h>

#include "gemm.cuh"
#include "kernels.cuh"

namespace ptts {

enum MegaPhaseType : int { MP_GEMM = 0, MP_ROWS = 1, MP_ATTN = 2 };
enum MegaRowOp : int { MR_EMBED = 0, MR_RES_LN = 1, MR_FINAL = 2, MR_SY = 3, MR_LNMOD = 4, MR_COMMIT = 5 };
enum MegaGemmMode : int { MG_PARTIAL = 0, MG_FULL_F32 = 1, MG_FULL_BF16 = 2 };

constexpr int MEGA_MAX_PHASES = 256;
constexpr int MEGA_SYNC_WORDS = MEGA_MAX_PHASES + 8;   // barrier counters, abort flag, exit counter
constexpr int MEGA_ABORT_WORD = MEGA_MAX_PHASES;
constexpr int MEGA_EXIT_WORD = MEGA_MAX_PHASES + 1;

struct MegaGemm {
  const float* bias;   // full modes (may be null)
  void* out;           // partial slabs [splits][rows_pad][nf] fp32, or the output matrix
  int wmap, amap;      // tensor-map indices
  int n_tile;          // features per item (UMMA N): 16, 32 or 64
  int tiles_n, splits, kb_per_split;
  int nf;              // output features
  int mode, act, out_ld;
};

struct alignas(16) MegaPhase {
  int type, n_items, n_gemm, op;
  int S;               // partial slabs summed by this phase's consumer (rows / attention)
  int d;               // row width
  int idx;             // layer (ATTN), LSD step (MR_SY, MR_COMMIT), block (MR_LNMOD)
  int flag;            // MR_SY: first LSD step; MR_COMMIT: last LSD step; MR_LNMOD: pending residual update
  int mod_ld;
  float eps;
  const float* part;   // [S][rows_pad][d]
  const float* bias;
  const float* w;      // LayerNorm affine (may be null)
  const float* b;
  const float* gate;   // AdaLN gate of the pending update (row stride mod_ld)
  const float* shift;
  const float* scale;
  void* out_act;       // bf16 operand written by the row op
  MegaGemm g[2];
};

struct MegaParams {
  const CUtensorMap* maps;
  const MegaPhase* phases;
  unsigned* sync;
  unsigned long long* trace;   // optional [G][n_phases + 1] %globaltimer stamps (PTTS_MEGA_TRACE=1), else null
  unsigned* diag;      // host-mapped: [0] code of a timed-out wait, [1] CTA, [2] thread (0 = healthy)
  int n_phases, n, G, stages, a_bytes, rows_pad, lsd;
  int D, H, F, Ld;
  const int* slot_ids;
  int* offsets;
  const float* rope_freqs;
  KvGeom kv;
  float *h, *cemb, *x_lat, *u, *eos_logit, *prev_latent, *commit_out;
  const float *noise, *bos, *w_in, *eos_w, *eos_b, *tcomb, *cond_b, *w_ip, *b_ip;
  bf16* attn_o;
};

size_t flow_mega_smem_bytes(int stages, int a_bytes);
cudaError_t flow_mega_launch(const MegaParams& P, cudaStream_t st);

}  // namespace ptts
