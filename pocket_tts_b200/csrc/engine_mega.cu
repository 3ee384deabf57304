// Engine implementation, part 4: the phase program of the persistent FlowLM step kernel
// (flow_mega.h) — tensor maps, phase list, launch.
#include <algorithm>
#include <cstdio>
#include <cstdlib>
#include <cstring>

#include "engine.h"

namespace ptts {

static const unsigned* g_mega_diag = nullptr;   // host view of the last engine's diagnostic words
const unsigned* mega_diag_words() { return g_mega_diag; }

bool Engine::mega_eligible(int n) const {
  const ptts_config& c = cfg;
  return use_mega && !batch_shared && kv_flow.type != 2 && act_type == 1 && use_tc && n <= mega_rows_pad && c.d_model == 1024 && c.flow_dim == 512 &&
         c.ldim == 32 && c.ffn % 64 == 0 && c.n_heads * 64 == c.d_model;
}

int Engine::mega_init() {
  if (const char* e = getenv("PTTS_MEGA")) use_mega = !(e[0] == '0');
  if (const char* e = getenv("PTTS_MEGA_CTAS")) mega_ctas = atoi(e);
  if (const char* e = getenv("PTTS_MEGA_STAGES")) mega_stages = atoi(e);
  mega_ctas = std::max(8, std::min(mega_ctas, n_sms));
  mega_rows_pad = cfg.max_streams <= 64 ? 64 : 128;
  if (cfg.max_streams > 128) use_mega = false;
  if (!use_mega || act_type != 1) return PTTS_OK;
  d_mega_sync = dmalloc<unsigned>(allocs, MEGA_SYNC_WORDS, stream);
  mega_part = dmalloc<float>(allocs, (size_t)8 * mega_rows_pad * std::max(cfg.d_model, 3 * cfg.d_model / 4), stream);
  if (!d_mega_sync || !mega_part) PTTS_FAIL(PTTS_ECUDA, "out of device memory for the step kernel's scratch");
  PTTS_CUDA_CHECK(cudaHostAlloc((void**)&h_mega_diag, 64, cudaHostAllocMapped));
  std::memset(h_mega_diag, 0, 64);
  PTTS_CUDA_CHECK(cudaHostGetDevicePointer((void**)&d_mega_diag, h_mega_diag, 0));
  g_mega_diag = h_mega_diag;
  return PTTS_OK;
}

int Engine::mega_plan(int lsd, MegaPlan** out) {
  auto it = mega_plans.find(lsd);
  if (it != mega_plans.end()) { *out = &it->second; return PTTS_OK; }
  const ptts_config& c = cfg;
  const int D = c.d_model, F = c.flow_dim, Ld = c.ldim, S = c.max_streams;
  const int modld = (c.flow_depth * 3 + 2) * F;
  std::vector<CUtensorMap> maps;
  std::vector<MegaPhase> ph;
  bool ok = true;
  auto add_map = [&](const void* base, long long rows, int K, int box_rows) -> int {
    CUtensorMap m;
    if (!tc_make_map(&m, base, rows, K, K, box_rows)) ok = false;
    maps.push_back(m);
    return (int)maps.size() - 1;
  };
  const int am_aln = add_map(a_ln, S, D, mega_rows_pad), am_ao = add_map(attn_o, S, D, mega_rows_pad);
  const int am_mid = add_map(mid, S, c.ffn, mega_rows_pad), am_cc = add_map(c_cond, S, D, mega_rows_pad);
  const int am_sy = add_map(sy, S, F, mega_rows_pad), am_g = add_map(g, S, F, mega_rows_pad);
  const int am_g1 = add_map(g1, S, F, mega_rows_pad);
  auto gemm = [&](const Mat& W, int amap, int n_tile, int splits, int mode, int act, const float* bias, void* out,
                  int out_ld) {
    MegaPhase p{};
    p.type = MP_GEMM; p.n_gemm = 1;
    MegaGemm& gg = p.g[0];
    const int kb = W.K / 64;
    if (!W.b16 || W.K % 64 || W.N % n_tile || kb % splits) ok = false;
    gg.bias = bias; gg.out = out; gg.wmap = add_map(W.b16, W.N, W.K, n_tile); gg.amap = amap;
    gg.n_tile = n_tile; gg.tiles_n = W.N / n_tile; gg.splits = splits; gg.kb_per_split = kb / splits;
    gg.nf = W.N; gg.mode = mode; gg.act = act; gg.out_ld = out_ld;
    p.n_items = gg.tiles_n * gg.splits;
    ph.push_back(p);
  };
  auto rows = [&](int op, int per_row) {
    MegaPhase p{};
    p.type = MP_ROWS; p.op = op; p.n_items = -per_row; p.part = mega_part; p.mod_ld = modld;
    return p;
  };
  constexpr int SK = 8;   // K splits of the GEMMs whose consumer is a row-wise phase
  { MegaPhase p = rows(MR_EMBED, D / 128); ph.push_back(p); }
  for (int l = 0; l < c.n_layers; ++l) {
    const LayerW& w = flow_layers[l];
    { MegaPhase p = rows(MR_RES_LN, 1); p.d = D; p.S = l == 0 ? 0 : SK; p.w = w.n1w; p.b = w.n1b; p.eps = 1e-5f; p.out_act = a_ln; ph.push_back(p); }
    gemm(w.in_proj, am_aln, 64, 2, MG_PARTIAL, ACT_NONE, nullptr, mega_part, 0);
    ph.back().flag = 1; ph.back().idx = l;   // compute warps prefetch this layer's cached K/V into L2 meanwhile
    { MegaPhase p{}; p.type = MP_ATTN; p.n_items = -c.n_heads; p.S = 2; p.part = mega_part; p.idx = l; ph.push_back(p); }
    gemm(w.out_proj, am_ao, 64, SK, MG_PARTIAL, ACT_NONE, nullptr, mega_part, 0);
    { MegaPhase p = rows(MR_RES_LN, 1); p.d = D; p.S = SK; p.w = w.n2w; p.b = w.n2b; p.eps = 1e-5f; p.out_act = a_ln; ph.push_back(p); }
    gemm(w.lin1, am_aln, 32, 1, MG_FULL_BF16, ACT_GELU, nullptr, mid, c.ffn);
    gemm(w.lin2, am_mid, 64, SK, MG_PARTIAL, ACT_NONE, nullptr, mega_part, 0);
  }
  { MegaPhase p = rows(MR_FINAL, 1); p.d = D; p.S = SK; p.w = out_norm_w; p.b = out_norm_b; p.eps = 1e-5f; p.out_act = c_cond; ph.push_back(p); }
  gemm(cond_embed, am_cc, 64, SK, MG_PARTIAL, ACT_NONE, nullptr, mega_part, 0);
  for (int s = 0; s < lsd; ++s) {
    { MegaPhase p = rows(MR_SY, 2); p.d = F; p.S = SK; p.idx = s; p.flag = s == 0; p.out_act = sy; ph.push_back(p); }
    gemm(mod_all, am_sy, 64, 1, MG_FULL_F32, ACT_NONE, mod_b, mod, modld);
    for (int j = 0; j <= c.flow_depth; ++j) {
      const bool last = j == c.flow_depth;
      MegaPhase p = rows(MR_LNMOD, 1);
      p.d = F; p.S = SK; p.idx = j; p.flag = j > 0; p.eps = 1e-6f; p.out_act = g;
      if (j > 0) { p.bias = flow_blocks[j - 1].b2; p.gate = mod + (size_t)(j - 1) * 3 * F + 2 * F; }
      if (!last) { p.w = flow_blocks[j].lnw; p.b = flow_blocks[j].lnb; }
      p.shift = mod + (size_t)j * 3 * F;   // block j: shift | scale | gate; final layer: shift | scale (mlp.py:108,128)
      p.scale = p.shift + F;
      ph.push_back(p);
      if (last) break;
      gemm(flow_blocks[j].mlp0, am_g, 16, 1, MG_FULL_BF16, ACT_SILU, flow_blocks[j].b0, g1, F);
      gemm(flow_blocks[j].mlp2, am_g1, 64, SK, MG_PARTIAL, ACT_NONE, nullptr, mega_part, 0);
    }
    gemm(final_lin, am_g, Ld, SK, MG_PARTIAL, ACT_NONE, nullptr, mega_part, 0);
    { MegaPhase p = rows(MR_COMMIT, 1); p.d = Ld; p.S = SK; p.bias = final_b; p.idx = s; p.flag = s == lsd - 1; ph.push_back(p); }
  }
  if (!ok) PTTS_FAIL(PTTS_ESTATE, "step kernel: a GEMM of this model does not fit the phase program (tile / split / tensor map)");
  if ((int)ph.size() > MEGA_MAX_PHASES) PTTS_FAIL(PTTS_ESTATE, "step kernel: %zu phases exceed %d", ph.size(), MEGA_MAX_PHASES);
  MegaPlan plan;
  plan.n_phases = (int)ph.size();
  PTTS_CUDA_CHECK(cudaMalloc((void**)&plan.d_maps, maps.size() * sizeof(CUtensorMap)));
  allocs.push_back(plan.d_maps);
  PTTS_CUDA_CHECK(cudaMalloc((void**)&plan.d_phases, ph.size() * sizeof(MegaPhase)));
  allocs.push_back(plan.d_phases);
  // plain synchronous copies: plans are built outside stream capture (see Engine::frame callers)
  PTTS_CUDA_CHECK(cudaMemcpy(plan.d_maps, maps.data(), maps.size() * sizeof(CUtensorMap), cudaMemcpyHostToDevice));
  PTTS_CUDA_CHECK(cudaMemcpy(plan.d_phases, ph.data(), ph.size() * sizeof(MegaPhase), cudaMemcpyHostToDevice));
  auto ins = mega_plans.emplace(lsd, plan);
  *out = &ins.first->second;
  return PTTS_OK;
}

int Engine::flow_sequence_mega(int n, int lsd, const float* noise, float* eos_out) {
  MegaPlan* plan = nullptr;
  int rc;
  if ((rc = mega_plan(lsd, &plan)) != PTTS_OK) return rc;
  const ptts_config& c = cfg;
  MegaParams P{};
  P.maps = plan->d_maps; P.phases = plan->d_phases; P.sync = d_mega_sync; P.diag = d_mega_diag;
  P.n_phases = plan->n_phases; P.n = n; P.G = mega_ctas; P.stages = mega_stages;
  P.rows_pad = mega_rows_pad; P.a_bytes = mega_rows_pad * 128; P.lsd = lsd;
  P.D = c.d_model; P.H = c.n_heads; P.F = c.flow_dim; P.Ld = c.ldim;
  P.slot_ids = d_slot_ids; P.offsets = offsets; P.rope_freqs = rope_freqs; P.kv = kv_flow;
  P.h = h; P.cemb = cemb; P.x_lat = x_lat; P.u = u; P.eos_logit = eos_out; P.prev_latent = prev_latent;
  P.commit_out = commit_out; P.noise = noise; P.bos = bos_emb; P.w_in = w_in; P.eos_w = eos_w; P.eos_b = eos_b;
  P.tcomb = tcomb; P.cond_b = cond_b; P.w_ip = input_proj.f32; P.b_ip = inproj_b;
  P.attn_o = reinterpret_cast<bf16*>(attn_o);
  static const bool trace_on = [] { const char* e = getenv("PTTS_MEGA_TRACE"); return e && e[0] == '1'; }();
  cudaStreamCaptureStatus cs = cudaStreamCaptureStatusNone;
  cudaStreamIsCapturing(stream, &cs);
  const bool tracing = trace_on && cs == cudaStreamCaptureStatusNone;
  const size_t tw = (size_t)mega_ctas * (plan->n_phases + 1);
  if (tracing) {
    if (!d_mega_trace) PTTS_CUDA_CHECK(cudaMalloc((void**)&d_mega_trace, (size_t)n_sms * (MEGA_MAX_PHASES + 1) * 8));
    PTTS_CUDA_CHECK(cudaMemsetAsync(d_mega_trace, 0, tw * 8, stream));
    P.trace = d_mega_trace;
  }
  prof_begin("flow.step_kernel");
  PTTS_CUDA_CHECK(flow_mega_launch(P, stream));
  prof_end();
  ++launches;
  if (tracing) {   // developer aid (eager mode only): per-phase time between grid barriers
    std::vector<unsigned long long> t(tw);
    PTTS_CUDA_CHECK(cudaStreamSynchronize(stream));
    PTTS_CUDA_CHECK(cudaMemcpy(t.data(), d_mega_trace, tw * 8, cudaMemcpyDeviceToHost));
    std::vector<MegaPhase> ph(plan->n_phases);
    PTTS_CUDA_CHECK(cudaMemcpy(ph.data(), plan->d_phases, ph.size() * sizeof(MegaPhase), cudaMemcpyDeviceToHost));
    const int np1 = plan->n_phases + 1;
    unsigned long long t0 = ~0ull, prev_end = 0;
    for (int b = 0; b < mega_ctas; ++b) t0 = std::min(t0, t[(size_t)b * np1]);
    prev_end = t0;
    static int dumps = 0;
    if (dumps++ == 3) {
      fprintf(stderr, "[step kernel trace] %d phases, %d CTAs\n", plan->n_phases, mega_ctas);
      for (int p = 0; p < plan->n_phases; ++p) {
        unsigned long long first = ~0ull, last = 0;
        for (int b = 0; b < mega_ctas; ++b) { const unsigned long long v = t[(size_t)b * np1 + p + 1]; first = std::min(first, v); last = std::max(last, v); }
        fprintf(stderr, "  phase %3d type %d op %d idx %2d items %4d : barrier-to-barrier %6.2f us (first CTA done after %6.2f us)  t=%7.2f\n", p,
                ph[p].type, ph[p].op, ph[p].idx, ph[p].n_items, (last - prev_end) * 1e-3, ((long long)first - (long long)prev_end) * 1e-3, (last - t0) * 1e-3);
        prev_end = last;
      }
    }
  }
  return PTTS_OK;
}

}  // namespace ptts
