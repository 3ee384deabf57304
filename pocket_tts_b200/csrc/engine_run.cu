// Engine implementation, part 3: the kernel sequences of prefill / frame step / Mimi decode.
#include "engine.h"

#include <algorithm>
#include <climits>
#include <cstdlib>
#include <cstring>

namespace ptts {

static const char* kTrTag[3] = {"seanet.convtr0", "seanet.convtr1", "seanet.convtr2"};
static const char* kC3Tag[3] = {"seanet.res0.conv3", "seanet.res1.conv3", "seanet.res2.conv3"};
static const char* kC1Tag[3] = {"seanet.res0.conv1", "seanet.res1.conv1", "seanet.res2.conv1"};
static inline RowMap plain(int ld) { return RowMap{0, 1 << 30, 1 << 30, ld}; }
static inline RowMap batched(long long bs, int rpb, int valid, int ld) { return RowMap{bs, rpb, valid, ld}; }

// developer timing probe: PTTS_SKIP_TAGS="flow.attn_decode,flow.ffn1" drops those launches (results are garbage;
// the change in step time is the in-graph cost of the kernel on the dependent chain)
static bool skip_tag(const char* tag) {
  static const char* list = getenv("PTTS_SKIP_TAGS");
  if (!list) return false;
  const size_t n = strlen(tag);
  for (const char* p = strstr(list, tag); p; p = strstr(p + 1, tag))
    if ((p == list || p[-1] == ',') && (p[n] == 0 || p[n] == ',')) return true;
  return false;
}

#define LAUNCH(tag, expr)       \
  do {                          \
    if (skip_tag(tag)) break;   \
    prof_begin(tag);            \
    PTTS_CUDA_CHECK(expr);      \
    prof_end();                 \
    ++launches;                 \
  } while (0)

// D = epilogue(A W^T).  a_is_act: A stored in the activation type (else fp32);
// out_is_act: out / out2 / res stored in the activation type (else fp32).
// K/V rows of FlowLM layer `layer` for the n streams in d_slot_ids: (k|v, head) regions of kv_capacity rows each
KvPrefetchJob Engine::flow_kv_prefetch_job(int layer, int n) const {
  KvPrefetchJob j;
  if (!use_kv_side_prefetch || kv_flow.ring) return j;
  const long long region = (long long)kv_flow.cap * kv_flow.D * (long long)act_size;
  j.regions = 2 * kv_flow.H;
  j.region_stride = region;
  j.slot_stride = region * j.regions;
  j.base = reinterpret_cast<const char*>(kv_flow.base) + (long long)layer * kv_flow.n_slots * j.slot_stride;
  j.slot_ids = d_slot_ids; j.offsets = offsets;
  j.row_bytes = kv_flow.D * (int)act_size;
  j.n = n;
  return j;
}

int Engine::gemm(const char* tag, const void* A, bool a_is_act, RowMap amap, const Mat& W, int M,
                 GemmEpilogue ep, bool out_is_act, const KvPrefetchJob* kv_prefetch) {
  GemmProblem p;
  if (kv_prefetch) p.kv_prefetch = *kv_prefetch;
  p.A = A; p.a_map = amap; p.M = M; p.N = W.N; p.K = W.K; p.ep = ep;
  {  // developer override of the split-K factor per tag: PTTS_SPLITS="flow.ffn1:2,flow.qkv:4"
    static const char* list = getenv("PTTS_SPLITS");
    if (list) {
      const size_t n = strlen(tag);
      for (const char* q = strstr(list, tag); q; q = strstr(q + 1, tag))
        if ((q == list || q[-1] == ',') && q[n] == ':') { p.max_splits = atoi(q + n + 1); break; }
    }
  }
  const int in_type = a_is_act ? act_type : 0;
  const int out_type = out_is_act ? act_type : 0;
  if (in_type == 1) {
    if (!W.b16) PTTS_FAIL(PTTS_ESTATE, "bf16 GEMM on a matrix without bf16 copy");
    p.W = W.b16;
  } else {
    p.W = W.f32;
  }
  if (in_type == 0 && out_type == 1) PTTS_FAIL(PTTS_ESTATE, "unsupported GEMM type combination");
  const bool plain_a = amap.rows_per_batch >= (1 << 30);
  const bool skinny = plain_a && M <= 256 && ep.out_map.rows_per_batch >= (1 << 30);
  if (in_type == 1 && use_tc && (p.K % 64) == 0 && (skinny || p.N % 32 == 0 || p.N == 16)) {
    // tcgen05 path: A rows addressable through the (possibly overlapping-row) tensor map
    const long long a_rows = plain_a ? M : (long long)M - (p.K / amap.ld - 1);
    LAUNCH(tag, gemm_tc_launch(p, out_type, a_rows, n_sms, pdl_enabled(), stream));
    return PTTS_OK;
  }
  LAUNCH(tag, gemm_simt_launch(p, in_type, out_type, stream));
  return PTTS_OK;
}

// pre-LN transformer layer on `rows` rows of the fp32 residual stream hbuf (in place):
// mimi_transformer.py:39-54.  The attention kernel is supplied by the caller via `mimi`.
int Engine::transformer_layer_common(const LayerW& L, float* hbuf, int rows, int d, int ffn, void* a,
                                     float* qkvb, void* ao, void* midb, bool mimi, int li, int n) {
  int rc;
  LAUNCH(mimi ? "mimi.ln" : "flow.ln", launch_ln_rows(hbuf, L.n1w, L.n1b, 1e-5f, a, act_type, rows, d, stream));
  {
    GemmEpilogue ep; ep.out = qkvb; ep.out_map = plain(3 * d);
    if ((rc = gemm(mimi ? "mimi.qkv" : "flow.qkv", a, true, plain(d), L.in_proj, rows, ep, false)) != PTTS_OK) return rc;
  }
  if (mimi) {
    LAUNCH("mimi.attn", launch_attn_window(qkvb, rope_freqs, mimi_offsets, d_slot_ids, kv_mimi, li, cfg.mimi_context,
                                           cfg.upsample_stride, ao, act_type, n, stream, rows / (n * cfg.upsample_stride), mimi_alone));
  } else if (rows == n) {
    LAUNCH("flow.attn_decode", launch_attn_decode(qkvb, rope_freqs, offsets, d_slot_ids, flow_kv(), li, ao, act_type, n, stream));
  } else {
    LAUNCH("flow.attn_prefill", launch_attn_chunk(qkvb, rope_freqs, offsets, d_slot_ids, d_row_start, d_lens, flow_kv(), li,
                             INT_MAX / 2, ao, act_type, n, stream, prefill_max_len));
  }
  {
    GemmEpilogue ep; ep.out = hbuf; ep.out_map = plain(d); ep.res = hbuf; ep.res_map = plain(d);
    ep.colscale = L.ls1;
    if ((rc = gemm(mimi ? "mimi.out_proj" : "flow.out_proj", ao, true, plain(d), L.out_proj, rows, ep, false)) != PTTS_OK) return rc;
  }
  LAUNCH(mimi ? "mimi.ln" : "flow.ln", launch_ln_rows(hbuf, L.n2w, L.n2b, 1e-5f, a, act_type, rows, d, stream));
  {
    GemmEpilogue ep; ep.out = midb; ep.out_map = plain(ffn); ep.act_pre = ACT_GELU;
    if ((rc = gemm(mimi ? "mimi.ffn1" : "flow.ffn1", a, true, plain(d), L.lin1, rows, ep, true)) != PTTS_OK) return rc;
  }
  {
    GemmEpilogue ep; ep.out = hbuf; ep.out_map = plain(d); ep.res = hbuf; ep.res_map = plain(d);
    ep.colscale = L.ls2;
    if ((rc = gemm(mimi ? "mimi.ffn2" : "flow.ffn2", midb, true, plain(ffn), L.lin2, rows, ep, false)) != PTTS_OK) return rc;
  }
  return PTTS_OK;
}

// FlowLM decode layer (one new position per stream, bf16 mode) with both LayerNorms folded into the GEMMs around
// them: the residual GEMMs (out_proj, linear2) also emit a bf16 copy of the new residual row and per-block
// (sum, M2) partials; the next GEMM (linear1 / the next layer's in_proj) runs on that copy with gamma folded into
// its weights and applies rstd * (acc - mean * c1) + c2 in its epilogue.  5 launches per layer instead of 7.
// Layer 0 still normalises the embedding rows with ln_rows (their producer is not a GEMM).
int Engine::flow_layer_decode_folded(int l, int n) {
  const LayerW& L = flow_layers[l];
  const int d = cfg.d_model, ffn = cfg.ffn, np = d / 16;
  const KvPrefetchJob pf = flow_kv_prefetch_job(l, n);
  int rc;
  if (l == 0) {
    LAUNCH("flow.ln", launch_ln_rows(h, L.n1w, L.n1b, 1e-5f, a_ln, act_type, n, d, stream));
    GemmEpilogue ep; ep.out = qkv; ep.out_map = plain(3 * d);
    if ((rc = gemm("flow.qkv", a_ln, true, plain(d), L.in_proj, n, ep, false, &pf)) != PTTS_OK) return rc;
  } else {
    GemmEpilogue ep; ep.out = qkv; ep.out_map = plain(3 * d);
    ep.ln_part_in = ln_part; ep.ln_nparts = np; ep.ln_c1 = L.qkv_c1; ep.ln_c2 = L.qkv_c2;
    if ((rc = gemm("flow.qkv", h16, true, plain(d), L.in_proj_ln, n, ep, false, &pf)) != PTTS_OK) return rc;
  }
  LAUNCH("flow.attn_decode", launch_attn_decode(qkv, rope_freqs, offsets, d_slot_ids, flow_kv(), l, attn_o, act_type, n, stream));
  {
    GemmEpilogue ep; ep.out = h; ep.out_map = plain(d); ep.res = h; ep.res_map = plain(d);
    ep.ln_raw = h16; ep.ln_part_out = ln_part;
    if ((rc = gemm("flow.out_proj", attn_o, true, plain(d), L.out_proj, n, ep, false)) != PTTS_OK) return rc;
  }
  {
    GemmEpilogue ep; ep.out = mid; ep.out_map = plain(ffn); ep.act_pre = ACT_GELU;
    static const bool no_gelu = getenv("PTTS_DEBUG_NOGELU") != nullptr;   // timing probe only (wrong results)
    if (no_gelu) ep.act_pre = ACT_NONE;
    ep.ln_part_in = ln_part; ep.ln_nparts = np; ep.ln_c1 = L.ffn_c1; ep.ln_c2 = L.ffn_c2;
    if ((rc = gemm("flow.ffn1", h16, true, plain(d), L.lin1_ln, n, ep, true)) != PTTS_OK) return rc;
  }
  {
    GemmEpilogue ep; ep.out = h; ep.out_map = plain(d); ep.res = h; ep.res_map = plain(d);
    if (l + 1 < cfg.n_layers) { ep.ln_raw = h16; ep.ln_part_out = ln_part; }
    if ((rc = gemm("flow.ffn2", mid, true, plain(ffn), L.lin2, n, ep, false)) != PTTS_OK) return rc;
  }
  return PTTS_OK;
}

// FlowLM step + flow head for the n streams in d_slot_ids; noise in noise_buf.
int Engine::flow_sequence(int n, int lsd, const float* noise, float* eos_out) {
  if (!noise) noise = noise_buf;
  if (!eos_out) eos_out = eos_logit;
  const ptts_config& c = cfg;
  const int D = c.d_model, F = c.flow_dim, Ld = c.ldim;
  const int modld = (c.flow_depth * 3 + 2) * F;
  int rc;
  if (mega_eligible(n)) return flow_sequence_mega(n, lsd, noise, eos_out);
  // u = input_proj(x_0) of the first Euler step depends on nothing but the noise: it runs on a side branch beside
  // the transformer layers instead of sitting in the dependent chain of the flow head
  // bf16 mode: everything after the AdaLN modulation GEMM of an Euler step is one cluster kernel (flow_head.cu)
  const bool head_cluster = head_mode != 0 && act_type == 1 && use_tc && flow_head_ss_supported(F, Ld, c.flow_depth);
  const bool side_u = use_side_branch && side_stream != nullptr && !head_cluster;
  if (side_u) {
    PTTS_CUDA_CHECK(cudaEventRecord(ev_side_fork, stream));
    PTTS_CUDA_CHECK(cudaStreamWaitEvent(side_stream, ev_side_fork, 0));
    cudaStream_t main_stream = stream;
    stream = side_stream;
    GemmEpilogue ep; ep.bias = inproj_b; ep.out = u; ep.out_map = plain(F);
    rc = gemm("head.input_proj", noise, false, plain(Ld), input_proj, n, ep, false);
    stream = main_stream;
    if (rc != PTTS_OK) return rc;
    PTTS_CUDA_CHECK(cudaEventRecord(ev_side_join, side_stream));
  }
  LAUNCH("flow.embed", launch_flow_embed(prev_latent, d_slot_ids, bos_emb, w_in, h, n, D, Ld, stream));
  static const int ln_fold_max = [] { const char* e = getenv("PTTS_LNFOLD_MAX"); return e ? std::min(atoi(e), 128) : 128; }();   // measured at 256 rows: +1.8 % for the fold, -4 % for the ring stage its statistics table costs
  const bool ln_fold = use_ln_fold && act_type == 1 && use_tc && n <= ln_fold_max && D % 64 == 0;
  for (int l = 0; l < c.n_layers; ++l)
    if ((rc = ln_fold ? flow_layer_decode_folded(l, n)
                      : transformer_layer_common(flow_layers[l], h, n, D, c.ffn, a_ln, qkv, attn_o, mid, false, l, n)) != PTTS_OK)
      return rc;
  // out_norm + EOS logit; the stream offsets advance here (every attention kernel of the step has read them)
  LAUNCH("flow.final_norm_eos", launch_final_norm_eos(h, out_norm_w, out_norm_b, 1e-5f, eos_w, eos_b, c_cond, act_type, eos_out, n, D, stream, offsets, d_slot_ids));
  // LSD / Euler integration from the injected noise (flow_lm.py:32-40).  x_0 = noise is read in place: the first
  // input_proj takes `noise` as its operand and the first Euler update writes x_lat = noise + v / lsd.
  const bool fused_sy = lsd == 1 && cond_b_t0 != nullptr;
  // where the gate opens: 0 = when the AdaLN modulation GEMM has finished (the head kernel starts), 2 = when it is
  // launched (cond_embed finished; the Mimi kernels arrive ~5-8 us after the event), 1 = before cond_embed (measured: worse,
  // the Mimi kernels take the SMs the 80-CTA AdaLN GEMM needs)
  static const int gate_early = [] { const char* e = getenv("PTTS_GATE_EARLY"); return e ? atoi(e) : 2; }();   // 2: 0.4635 -> 0.4594 ms against 0
  bool gate_joined = true;
  auto open_gate = [&]() -> int {
    PTTS_CUDA_CHECK(cudaEventRecord(ev_side_fork, stream));
    PTTS_CUDA_CHECK(cudaStreamWaitEvent(side_stream, ev_side_fork, 0));
    cudaStreamCaptureStatus cs = cudaStreamCaptureStatusNone;
    PTTS_CUDA_CHECK(cudaStreamIsCapturing(stream, &cs));
    if (cs == cudaStreamCaptureStatusActive) PTTS_CUDA_CHECK(cudaEventRecordWithFlags(gate_record, side_stream, cudaEventRecordExternal));
    else PTTS_CUDA_CHECK(cudaEventRecord(gate_record, side_stream));
    PTTS_CUDA_CHECK(cudaEventRecord(ev_side_join, side_stream));
    gate_joined = false;
    return PTTS_OK;
  };
  const bool gating = gate_record != nullptr && side_stream != nullptr && head_cluster;
  if (gate_early == 1 && gating) {
    if ((rc = open_gate()) != PTTS_OK) return rc;
  }
  if (fused_sy) {   // sy = SiLU(cond_embed(c) + b + (te(s)+te(t))/2) straight out of the GEMM epilogue (mlp.py:203-208)
    GemmEpilogue ep; ep.bias = cond_b_t0; ep.act_pre = ACT_SILU; ep.out = sy; ep.out_map = plain(F);
    if ((rc = gemm("head.cond_embed", c_cond, true, plain(D), cond_embed, n, ep, true)) != PTTS_OK) return rc;
  } else {
    GemmEpilogue ep; ep.bias = cond_b; ep.out = cemb; ep.out_map = plain(F);
    if ((rc = gemm("head.cond_embed", c_cond, true, plain(D), cond_embed, n, ep, false)) != PTTS_OK) return rc;
  }
  for (int s = 0; s < lsd; ++s) {
    const float* x_in = s == 0 ? noise : x_lat;
    if (!fused_sy) LAUNCH("head.silu_add", launch_silu_add(cemb, tcomb + (size_t)s * F, sy, act_type, n, F, stream));
    if (gate_early == 2 && gating && s == 0 && (rc = open_gate()) != PTTS_OK) return rc;
    {
      GemmEpilogue ep; ep.bias = mod_b; ep.out = mod; ep.out_map = plain(modld);
      if ((rc = gemm("head.adaln_mod", sy, true, plain(F), mod_all, n, ep, false)) != PTTS_OK) return rc;
    }
    if (head_cluster) {
      FlowHeadParams hp{};
      for (int j = 0; j < c.flow_depth; ++j) {
        const FlowBlockW& w = flow_blocks[j];
        hp.blk[j] = FlowHeadBlock{w.mlp0.b16, w.mlp2.b16, w.b0, w.b2, w.lnw, w.lnb};
      }
      hp.depth = c.flow_depth; hp.n = n; hp.ldim = Ld; hp.mod_ld = modld;
      hp.x_in = x_in; hp.w_in = input_proj.f32; hp.b_in = inproj_b; hp.mod = mod;
      hp.w_final = final_lin.b16; hp.b_final = final_b; hp.alpha = 1.0f / (float)lsd; hp.x_out = x_lat;
      hp.slot_ids = d_slot_ids;
      const bool last = s == lsd - 1;
      hp.prev_latent = last ? prev_latent : nullptr;
      hp.commit_out = last ? commit_out : nullptr;
      // head-gated Mimi passes: "the flow head of this frame starts" as an external event record node on a side branch
      // that hangs off the AdaLN modulation GEMM (the main branch keeps its programmatic edge to the head kernel)
      const bool gate = gating && s == 0 && gate_early == 0;
      if (gate && (rc = open_gate()) != PTTS_OK) return rc;
      auto join_gate = [&]() -> int {
        if (!gate_joined) { PTTS_CUDA_CHECK(cudaStreamWaitEvent(stream, ev_side_join, 0)); gate_joined = true; }
        return PTTS_OK;
      };
      if (head_mode == 3 && flow_head_ss_supported(F, Ld, c.flow_depth)) {
        FlowHeadTcParams tp;
        tp.h = hp;
        if (!flow_head_ss_prepare(&tp)) PTTS_FAIL(PTTS_ECUDA, "tensor map encode failed (flow head)");
        LAUNCH("head.cluster", launch_flow_head_ss(tp, stream));
        if ((rc = join_gate()) != PTTS_OK) return rc;
      } else {
        FlowHeadTcParams tp;
        tp.h = hp;
        tp.xchg = head_l2_exchange ? head_xchg : nullptr;
        if (!flow_head_tc_prepare(&tp)) PTTS_FAIL(PTTS_ECUDA, "tensor map encode failed (flow head)");
        LAUNCH("head.cluster", launch_flow_head_tc(tp, stream));
        if ((rc = join_gate()) != PTTS_OK) return rc;
      }
      continue;
    }
    if (s == 0 && side_u) {
      PTTS_CUDA_CHECK(cudaStreamWaitEvent(stream, ev_side_join, 0));
    } else {
      GemmEpilogue ep; ep.bias = inproj_b; ep.out = u; ep.out_map = plain(F);
      if ((rc = gemm("head.input_proj", x_in, false, plain(Ld), input_proj, n, ep, false)) != PTTS_OK) return rc;
    }
    for (int j = 0; j < c.flow_depth; ++j) {
      const FlowBlockW& w = flow_blocks[j];
      const float* mj = mod + (size_t)j * 3 * F;  // shift | scale | gate  (mlp.py:108)
      LAUNCH("head.ln_modulate", launch_ln_modulate(u, w.lnw, w.lnb, mj, mj + F, modld, g, act_type, n, F, stream));
      {
        GemmEpilogue ep; ep.bias = w.b0; ep.act_pre = ACT_SILU; ep.out = g1; ep.out_map = plain(F);
        if ((rc = gemm("head.mlp0", g, true, plain(F), w.mlp0, n, ep, true)) != PTTS_OK) return rc;
      }
      {
        GemmEpilogue ep; ep.bias = w.b2; ep.gate = mj + 2 * F; ep.gate_ld = modld;
        ep.res = u; ep.res_map = plain(F); ep.out = u; ep.out_map = plain(F);
        if ((rc = gemm("head.mlp2", g1, true, plain(F), w.mlp2, n, ep, false)) != PTTS_OK) return rc;
      }
    }
    const float* mf = mod + (size_t)c.flow_depth * 3 * F;  // shift | scale  (mlp.py:128)
    LAUNCH("head.ln_modulate", launch_ln_modulate(u, nullptr, nullptr, mf, mf + F, modld, g, act_type, n, F, stream));
    {
      GemmEpilogue ep; ep.bias = final_b; ep.alpha = 1.0f / (float)lsd;
      ep.res = x_in; ep.res_map = plain(Ld); ep.out = x_lat; ep.out_map = plain(Ld);
      if ((rc = gemm("head.final", g, true, plain(F), final_lin, n, ep, false)) != PTTS_OK) return rc;
    }
  }
  if (!head_cluster) LAUNCH("head.commit", launch_commit_latent(x_lat, d_slot_ids, prev_latent, commit_out, n, Ld, stream));
  return PTTS_OK;
}

// Mimi decode of `frames` consecutive latents per stream, latents [frames][n, ldim] (device fp32) ->
// pcm_out [n, frames * 1920].  One pass over frames * 16 tokens per stream is arithmetically the same as
// `frames` single-frame passes (causal attention, causal convolutions with carried history).
// Segments (seg_lo..seg_hi, default all; the head-gated loop captures them as separate graphs): 0 = upsample +
// transformer + conv0, 1 = stage 0, 2 = stage 1, 3 = first half of the streams through the fused last stage,
// 4 = second half + history write-back.  A whole pass runs the fused last stage as one launch.
int Engine::mimi_sequence(int n, const float* latents, int frames, float* pcm_out, int seg_lo, int seg_hi) {
  auto in = [&](int seg) { return seg >= seg_lo && seg <= seg_hi; };
  const bool whole = seg_lo <= 0 && seg_hi >= MIMI_SEGS - 1;
  const ptts_config& c = cfg;
  const int Md = c.mimi_d, U = c.upsample_stride * frames;
  int rc;
  if (frames < 1 || frames > mimi_gmax) PTTS_FAIL(PTTS_EINVAL, "mimi pass of %d frames (max %d)", frames, mimi_gmax);
  if (!pcm_out) pcm_out = pcm;
  HistTable tab = hist_tab;
  for (int k = 0; k < tab.count; ++k) tab.d[k].T *= frames;
  if (in(0)) {
  {  // developer probe (PTTS_MIMI_PAD_NODES=k): k extra trivial kernel nodes per Mimi pass - measures what a graph node of the
     // Mimi branch costs the FlowLM chain through the shared launch front end
    static const int pad = [] { const char* e = getenv("PTTS_MIMI_PAD_NODES"); return e ? atoi(e) : 0; }();
    for (int k = 0; k < pad; ++k) LAUNCH("advance", launch_advance(mimi_offsets, d_slot_ids, nullptr, 0, n, stream));
  }
  LAUNCH("mimi.embed_upsample", launch_mimi_embed(latents, d_slot_ids, emb_std, emb_mean, wq, wup, e_prev, xm, n, Md, c.ldim, c.upsample_stride, frames, stream));
  for (int l = 0; l < c.mimi_layers; ++l)
    if ((rc = transformer_layer_common(mimi_layers[l], xm, n * U, Md, c.mimi_ffn, m_ln, mqkv, m_attn, m_mid, true, l, n)) != PTTS_OK)
      return rc;
  LAUNCH("advance", launch_advance(mimi_offsets, d_slot_ids, nullptr, U, n, stream));
  LAUNCH("seanet.hist_gather", launch_hist_gather(tab, d_slot_ids, act_type, n, stream));
  }
  const int h0 = c.kernel - 1;
  if (in(0)) LAUNCH("seanet.stage", launch_stage_rows(xm, b_conv0, act_type, n, U, Md, h0, stream));
  const size_t es = act_size;
  auto off = [&](void* base, long long elems) { return (void*)((char*)base + elems * (long long)es); };
  int T = U, C = Md;
  if (in(0)) {  // Conv1d 512->512 K7 (seanet.py:141-143) -> ELU -> input of the first transposed conv
    const int rpb = h0 + T;
    GemmEpilogue ep; ep.bias = conv0_b; ep.act_post = ACT_ELU;
    ep.out = off(b_tr[0], 1LL * C); ep.out_map = batched((long long)(1 + T) * C, rpb, T, C);
    if ((rc = gemm("seanet.conv0", b_conv0, true, batched((long long)rpb * C, rpb, T, C), conv0, n * rpb, ep, true)) != PTTS_OK) return rc;
  }
  for (int i = 0; i < 3; ++i) {
    const int r = c.ratios[i], T2 = T * r, C2 = C / 2, hc = hid_width(C2 / 2), rk = c.res_kernel - 1;
    if (i == 2 && use_fused_stage2 && act_type == 1 && use_tc &&
        seanet_stage2_fused_supported(C, C2, r, C2 / 2, hc, c.res_kernel, c.last_kernel)) {
      // the whole last stage + final conv in one persistent kernel (seanet_fused.cu): reads b_tr[2], writes PCM
      const int cap = fused_sms > 0 ? std::min(fused_sms, n_sms) : (fused_sms < 0 ? 1 << 30 : n_sms);
      const int n_a = whole ? n : n / 2;   // streams of segment 3 (a whole pass: one launch over all streams)
      for (int half = 0; half < 2; ++half) {
        const int s0 = half == 0 ? 0 : n_a, ns = half == 0 ? n_a : n - n_a;
        if (!in(3 + half) || ns <= 0) continue;
        const long long row_in = (long long)s0 * (1 + T) * C, row_hist = (long long)s0 * (2 + 4 * T) * C2;
        LAUNCH("seanet.stage2_fused",
               launch_seanet_stage2_fused((const bf16*)b_tr[2] + row_in, stages[2].tr.b16, stages[2].c3.b16, stages[2].c1.b16, fused2_consts,
                                          (bf16*)b_res[2] + row_hist, (bf16*)b_last + row_hist, pcm_out + (long long)s0 * 4 * T, ns, T,
                                          cap, pdl_enabled(), stream));
      }
      if (in(4)) LAUNCH("seanet.hist_scatter", launch_hist_scatter(tab, d_slot_ids, act_type, n, stream));
      return PTTS_OK;
    }
    if (!(i < 2 ? in(1 + i) : (in(3) || in(4)))) { T = T2; C = C2; continue; }
    if (i == 2 && !whole) PTTS_FAIL(PTTS_ESTATE, "segmented Mimi passes need the fused last stage");
    {  // ConvTranspose1d K=2r (seanet.py:147-152): raw -> b_raw (residual), ELU(raw) -> b_res
      const int rpb = 1 + T;
      GemmEpilogue ep; ep.bias = stages[i].tr_b; ep.act_post = ACT_ELU;
      ep.out2 = b_raw[i]; ep.out2_map = batched((long long)T2 * C2, rpb, T, r * C2);
      ep.out = off(b_res[i], (long long)rk * C2); ep.out_map = batched((long long)(rk + T2) * C2, rpb, T, r * C2);
      if ((rc = gemm(kTrTag[i], b_tr[i], true, batched((long long)rpb * C, rpb, T, C), stages[i].tr, n * rpb, ep, true)) != PTTS_OK) return rc;
    }
    if (i < 2 && use_fused_resblock && act_type == 1 && use_tc && seanet_resblock_fused_supported(C2, C2 / 2, hc, c.res_kernel)) {
      // conv3 -> ELU -> conv1 + skip -> ELU in one persistent kernel, hidden tensor on chip (seanet_fused.cu)
      LAUNCH("seanet.resblock_fused",
             launch_seanet_resblock_fused(C2, (const bf16*)b_res[i], (const bf16*)b_raw[i], stages[i].c3.b16, stages[i].c1.b16,
                                          res_b3_host[i].data(), res_b1_host[i].data(), (bf16*)b_tr[i + 1], 1, n, T2, n_sms, pdl_enabled(), stream));
      T = T2; C = C2;
      continue;
    }
    {  // ResBlock conv K3 C2 -> C2/2 on ELU(x), ELU on the way out (seanet.py:27-41)
      const int rpb = rk + T2;
      GemmEpilogue ep; ep.bias = stages[i].c3_b; ep.act_post = ACT_ELU;
      ep.out = b_hid[i]; ep.out_map = batched((long long)T2 * hc, rpb, T2, hc);
      if ((rc = gemm(kC3Tag[i], b_res[i], true, batched((long long)rpb * C2, rpb, T2, C2), stages[i].c3, n * rpb, ep, true)) != PTTS_OK) return rc;
    }
    {  // ResBlock conv K1 + skip, then the ELU that precedes the next layer
      const int nh = (i < 2) ? 1 : c.last_kernel - 1;
      void* nxt = (i < 2) ? b_tr[i + 1] : b_last;
      GemmEpilogue ep; ep.bias = stages[i].c1_b; ep.act_post = ACT_ELU;
      ep.res = b_raw[i]; ep.res_map = batched((long long)T2 * C2, T2, T2, C2);
      ep.out = off(nxt, (long long)nh * C2); ep.out_map = batched((long long)(nh + T2) * C2, T2, T2, C2);
      if ((rc = gemm(kC1Tag[i], b_hid[i], true, batched((long long)T2 * hc, T2, T2, hc), stages[i].c1, n * T2, ep, true)) != PTTS_OK) return rc;
    }
    T = T2; C = C2;
  }
  if (!whole) return PTTS_OK;   // (unfused last stage: whole passes only, checked above)
  LAUNCH("seanet.last_conv", launch_last_conv(b_last, act_type, last_w, last_b, pcm_out, n, T, C, c.last_kernel, stream));
  LAUNCH("seanet.hist_scatter", launch_hist_scatter(tab, d_slot_ids, act_type, n, stream));
  return PTTS_OK;
}

}  // namespace ptts
