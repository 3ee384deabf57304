// Launchers for the non-GEMM kernels of the frame step.  Element types are passed as
// 0 = float, 1 = bf16 ("act" = activation storage feeding GEMMs, "kv" = KV-cache storage).
#pragma once
#include <cuda.h>

#include "common.cuh"

namespace ptts {

// KV cache geometry: element (layer, slot, kv, h, pos, d) lives at
//   base + ((((layer*n_slots + slot)*2 + kv)*H + h)*cap + pos)*D + d
struct KvGeom {
  void* base;
  int n_slots, H, cap, D;
  int type;        // 0 float, 1 bf16, 2 fp8 e4m3 with one fp32 scale per cached row (`scale`, FlowLM cache only)
  int ring;        // 1: pos stored at pos & (cap-1) (Mimi sliding window), 0: linear
  int prefetch;    // decode attention form (PTTS_ATTN): 0 per-lane loads, 1 + L2 prefetch before the producer wait, 2 bulk-copy kernel
  // Shared voice prefix (ptts_stream_fork_shared): positions [0, pfx_len[slot]) of a slot with pfx_slot[slot] >= 0 are
  // read from that slot's cache at the same position index (one resident copy of the voice K/V for every stream forked
  // from it); nullptr: no stream of the launch shares a prefix.
  const int* pfx_slot = nullptr;
  const int* pfx_len = nullptr;
  float* scale = nullptr;   // type 2: scale of row (layer, slot, kv, h, pos) at the row's index (same layout without D)
};

// h[i,:] = input_linear(where(isnan(prev), bos, prev))            flow_lm.py:121-122
cudaError_t launch_flow_embed(const float* prev_latent, const int* slot_ids, const float* bos,
                              const float* w_in /*[d,ldim]*/, float* h, int n, int d, int ldim,
                              cudaStream_t st);

// h[r,:] = table[tokens[r],:]   LUT conditioner embedding (conditioners/text.py:74-76); d % 4 == 0
cudaError_t launch_embed_gather(const int* tokens, const float* table, float* h, int rows, int d, cudaStream_t st);

// y = LayerNorm(x; w, b, eps) rows of width d (<= 1024), x fp32 -> y (act type)
cudaError_t launch_ln_rows(const float* x, const float* w, const float* b, float eps, void* y,
                           int y_type, int rows, int d, cudaStream_t st);

// FlowLM decode attention for one new position per stream: RoPE(q,k) at offset[slot], append
// K/V, softmax(q K^T / sqrt(D)) V over positions [0, offset].   transformer.py:135-158
cudaError_t launch_attn_decode(const float* qkv /*[n,3*H*D]*/, const float* rope_freqs,
                               const int* offsets, const int* slot_ids, KvGeom kv, int layer,
                               void* out, int out_type, int n, cudaStream_t st);

// Multi-query variant (prefill, Mimi): stream i owns rows [row_start[i], row_start[i]+len[i])
// of qkv, appended at positions offset[slot]..; causal with optional window `context`.
// q is rotated in place inside qkv.
cudaError_t launch_attn_chunk(float* qkv, const float* rope_freqs, const int* offsets,
                              const int* slot_ids, const int* row_start, const int* lens, KvGeom kv,
                              int layer, int context, void* out, int out_type, int n, cudaStream_t st,
                              int max_len = 0 /* longest chunk of the batch (host-known); > 8 selects the two-pass form */);

// Mimi decoder attention: exactly T (<= 16) new tokens per stream (rows i*T.. of qkv), sliding
// window `context` over the ring cache; the visible window is staged once in shared memory.
cudaError_t launch_attn_window(const float* qkv, const float* rope_freqs, const int* offsets,
                               const int* slot_ids, KvGeom kv, int layer, int context, int T, void* out,
                               int out_type, int n, cudaStream_t st, int frames = 1, bool alone = false);

// offsets[slot_ids[i]] += inc[i] (inc == nullptr: += const_inc)
cudaError_t launch_advance(int* offsets, const int* slot_ids, const int* inc, int const_inc, int n,
                           cudaStream_t st);

// c = out_norm(h) (act type) ; eos_logit = w_eos . c + b_eos (from the fp32 normalised row)
// flow_lm.py:128-129,153-154
cudaError_t launch_final_norm_eos(const float* h, const float* w, const float* b, float eps,
                                  const float* w_eos, const float* b_eos, void* c, int c_type,
                                  float* eos_logit, int n, int d, cudaStream_t st, int* offsets = nullptr,
                                  const int* slot_ids = nullptr);

// sy = silu(cemb + tcomb_row) (mlp.py:209, 99-100 SiLU of adaLN_modulation)
cudaError_t launch_silu_add(const float* cemb, const float* tcomb_row, void* sy, int sy_type,
                            int n, int d, cudaStream_t st);

// g = LN_flow(u; w,b (nullable), eps=1e-6) * (1 + scale) + shift     mlp.py:16-17,49-55,107-109
cudaError_t launch_ln_modulate(const float* u, const float* w, const float* b, const float* shift,
                               const float* scale, int mod_ld, void* g, int g_type, int n, int d,
                               cudaStream_t st);

// One Euler step of the flow head after the AdaLN modulation GEMM, as a single kernel (flow_head_ss.cu / flow_head_tc.cu,
// bf16 mode, flow_dim 512, ldim 32): input_proj, the ResBlocks, the final layer and the latent commit.
#define PTTS_HEAD_MAX_DEPTH 8
struct FlowHeadBlock {
  const bf16 *w0, *w2;                   // mlp.0 / mlp.2 weights [512, 512] bf16
  const float *b0, *b2, *lnw, *lnb;
};
struct FlowHeadParams {
  FlowHeadBlock blk[PTTS_HEAD_MAX_DEPTH];
  int depth, n, ldim, mod_ld;
  const float* x_in;                     // [n, ldim] fp32: x of this Euler step
  const float* w_in;                     // input_proj weight [512, ldim] fp32
  const float* b_in;
  const float* mod;                      // [n, mod_ld] fp32: (shift | scale | gate) x depth, then shift | scale
  const bf16* w_final;                   // [ldim, 512] bf16
  const float* b_final;
  float alpha;                           // 1 / lsd_steps
  float* x_out;                          // [n, ldim]: x + alpha * v
  const int* slot_ids;                   // commit (last Euler step): prev_latent[slot] = x_out row, optional copy
  float* prev_latent;                    // nullable
  float* commit_out;                     // nullable
};
// tcgen05 form (flow_head_tc.cu): the same step with TMEM accumulators, a TMA weight ring and bulk DSMEM operand
// exchange.  `maps`: one SWIZZLE_128B tensor map per matrix (mlp.0 / mlp.2 of every block, then the final layer).
struct FlowHeadTcParams {
  FlowHeadParams h;
  alignas(64) CUtensorMap maps[2 * PTTS_HEAD_MAX_DEPTH + 1];
  alignas(64) CUtensorMap xmap;          // operand exchange scratch as [blocks * 64 rows, 64] bf16
  bf16* xchg = nullptr;                  // [clusters][2 buffers][8 blocks][64 rows][64] bf16 global scratch (nullptr: exchange through DSMEM)
  int trace = 0;                         // developer timing stamps (PTTS_HEAD_TRACE=1)
};
bool flow_head_tc_prepare(FlowHeadTcParams* p);   // fills `maps` from h (cached driver encodes)
cudaError_t launch_flow_head_tc(const FlowHeadTcParams& p, cudaStream_t st);
// stream-split form (flow_head_ss.cu, default): 16 streams per CTA, features on the TMEM lanes, no cross-CTA exchange;
// uses `maps` with 128-row weight boxes (its own prepare) and ignores xmap / xchg.
bool flow_head_ss_supported(int flow_dim, int ldim, int depth);
bool flow_head_ss_prepare(FlowHeadTcParams* p);
cudaError_t launch_flow_head_ss(const FlowHeadTcParams& p, cudaStream_t st);

// x[i,:] -> prev_latent[slot] and out (nullable)
cudaError_t launch_commit_latent(const float* x, const int* slot_ids, float* prev_latent,
                                 float* out, int n, int ldim, cudaStream_t st);

// tts_model.py:449-451 + resample.py:41-48: de-normalise, 1x1 projection, depthwise x16
// transposed conv with the previous frame's embedding as streaming state.
cudaError_t launch_mimi_embed(const float* latents, const int* slot_ids, const float* emb_std,
                              const float* emb_mean, const float* w_q /*[C,ldim]*/,
                              const float* w_up /*[C,2S]*/, float* e_prev /*[slots,C]*/,
                              float* xm /*[n*frames*S,C]*/, int n, int C, int ldim, int S, int frames,
                              cudaStream_t st);

// rows of xm (fp32 [n*T, C]) -> buf[i][hist + t][c] (act type)
cudaError_t launch_stage_rows(const float* xm, void* buf, int buf_type, int n, int T, int C,
                              int hist, cudaStream_t st);

// streaming-conv history: per conv input buffer, `hist` leading rows come from the per-slot
// state (gather) and the trailing `hist` rows of this frame go back to it (scatter).
struct HistDesc {
  void* buf;          // [n][hist+T][C] act type
  void* state;        // [slots][hist][C] act type
  int hist, T, C;
};
#define PTTS_MAX_HIST 8
struct HistTable {
  HistDesc d[PTTS_MAX_HIST];
  int count;
};
cudaError_t launch_hist_gather(const HistTable& t, const int* slot_ids, int type, int n, cudaStream_t st);
cudaError_t launch_hist_scatter(const HistTable& t, const int* slot_ids, int type, int n, cudaStream_t st);

// final Conv1d C->1, kernel K: pcm[i][t] = b + sum_k sum_c w[k*C + c] * buf[i][t+k][c]
cudaError_t launch_last_conv(const void* buf, int buf_type, const float* w, const float* b,
                             float* pcm, int n, int T, int C, int K, cudaStream_t st);

// Last SEANet decoder stage + final conv as one persistent tcgen05 kernel (seanet_fused.cu; bf16 mode, english / 24L shapes):
// x_buf [n][1 + T][128] -> pcm [n][4T]; conv histories in / out through the history rows of res_buf / last_buf.
bool seanet_stage2_fused_supported(int c_in, int c_out, int ratio, int hid, int hid_padded, int res_kernel, int last_kernel);
struct FusedStage2Consts {       // host copies of the stage's per-channel vectors (they travel as kernel parameters)
  float b_tr[256], b_c1[64], b_c3[32], w_last[192], b_last;
};
// grid_cap: at most that many CTAs (>= number of tiles: one tile per CTA, short CTAs that high-priority kernels of the
// other branch can slip between; number of SMs: persistent form)
cudaError_t launch_seanet_stage2_fused(const bf16* x_buf, const bf16* w_tr, const bf16* w3, const bf16* w1, const FusedStage2Consts& cst,
                                       bf16* res_buf, bf16* last_buf, float* pcm, int n, int T, int grid_cap, bool pdl, cudaStream_t st,
                                       unsigned long long* trace = nullptr);

// SEANet ResBlock of stages 0 / 1 as one persistent tcgen05 kernel (conv3 -> ELU -> conv1 + skip -> ELU, hidden tensor kept
// on chip): e_buf [n][2 + T][C] = ELU(y) behind two history rows, raw [n][T][C] = y, out [n][out_hist + T][C]; C = 128 | 256
bool seanet_resblock_fused_supported(int C, int hid, int hid_padded, int res_kernel);
cudaError_t launch_seanet_resblock_fused(int C, const bf16* e_buf, const bf16* raw, const bf16* w3, const bf16* w1, const float* b3_host,
                                         const float* b1_host, bf16* out, int out_hist, int n, int T, int n_sms, bool pdl, cudaStream_t st);

// data/audio.py:77-78: (clamp(x, -1, 1) * 32767) truncated to int16
cudaError_t launch_pcm_to_int16(const float* pcm, short* out, long long n, cudaStream_t st);

// fp32 -> bf16 conversion (weights at load time), strided fill helpers
cudaError_t launch_f32_to_bf16(const float* src, bf16* dst, long long n, cudaStream_t st);
cudaError_t launch_fill_f32(float* dst, float v, long long n, cudaStream_t st);
cudaError_t launch_bf16_to_f32(const bf16* src, float* dst, long long n, cudaStream_t st);

// KV import/export between the reference layout cache[2,1,P,H,D] (fp32) and the engine layout
cudaError_t launch_kv_import(const float* ref /*[2,P,H,D]*/, KvGeom kv, int layer, int slot, int P,
                             cudaStream_t st);
cudaError_t launch_kv_export(float* ref, KvGeom kv, int layer, int slot, int P, cudaStream_t st);
cudaError_t launch_kv_fork(KvGeom kv, int n_layers, int src, int dst, int P, cudaStream_t st);
cudaError_t launch_kv_export_range(float* ref, KvGeom kv, int layer, int slot, int P, int p0, int p1, cudaStream_t st);
cudaError_t launch_kv_fork_range(KvGeom kv, int n_layers, int src, int dst, int p0, int p1, cudaStream_t st);

}  // namespace ptts
