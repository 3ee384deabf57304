// Engine implementation, part 1: memory, weight loading and re-layout.
#include "engine.h"

#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>

namespace ptts {

static thread_local std::string g_err;
void set_error(const std::string& msg) { g_err = msg; }
void set_error_cuda(cudaError_t e, const char* expr, const char* file, int line) {
  char buf[512];
  snprintf(buf, sizeof buf, "CUDA error %s (%s) at %s:%d: %s", cudaGetErrorName(e), cudaGetErrorString(e),
           file, line, expr);
  g_err = buf;
  if (const unsigned* d = mega_diag_words())
    if (d[0]) {
      snprintf(buf, sizeof buf, " [step kernel: wait 0x%x timed out in CTA %u thread %u]", d[0], d[1], d[2]);
      g_err += buf;
    }
}
const char* get_error() { return g_err.c_str(); }
void prefer_max_smem_carveout(const void* kern) {
  static std::map<const void*, int> done;
  static int on = -1;
  if (on < 0) { const char* e = getenv("PTTS_CARVEOUT"); on = (e && e[0] == '1') ? 1 : 0; }
  if (!on || done.count(kern)) return;
  cudaFuncSetAttribute(kern, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared);
  done[kern] = 1;
}
bool pdl_enabled() {
  static int v = -1;
  if (v < 0) { const char* e = getenv("PTTS_PDL"); v = (e && e[0] == '0') ? 0 : 1; }
  return v == 1;
}

static void tl_dump(unsigned long long* dev) {
  std::vector<unsigned long long> b(2 + TL_W * TL_MAX);
  cudaDeviceSynchronize();
  if (cudaMemcpy(b.data(), dev, b.size() * 8, cudaMemcpyDeviceToHost) != cudaSuccess) return;
  const unsigned long long total = b[0];
  const unsigned n = (unsigned)std::min<unsigned long long>(total, 160);
  fprintf(stderr, "[timeline] %llu records, last %u (us relative to the first shown: entry / wait returned / exit of CTA 0)\n", total, n);
  unsigned long long t0 = 0;
  const unsigned long long last_attn_exit = b[1];
  for (unsigned k = 0; k < n; ++k) {
    const unsigned s = (unsigned)((total - n + k) % TL_MAX);
    const unsigned long long* r = &b[2 + TL_W * s];
    const unsigned long long kind = r[0], te = r[1], tw = r[2], tx = r[3];
    if (k == 0) t0 = te;
    char name[64];
    if (kind >> 28) snprintf(name, sizeof name, "gemm N=%llu kb=%llu", (kind >> 8) & 0xFFFFF, kind & 0xFF);
    else snprintf(name, sizeof name, "%s", kind == 2 ? "attn_decode" : kind == 3 ? "ln_rows" : kind == 4 ? "final_norm_eos" : kind == 5 ? "flow_embed" : kind == 6 ? "head_cluster" : "?");
    fprintf(stderr, "[timeline] %-22s entry %9.2f  wait %9.2f  exit %9.2f |", name, (double)(te - t0) / 1e3,
            tw ? (double)(tw - t0) / 1e3 : -1.0, tx ? (double)(tx - t0) / 1e3 : -1.0);
    for (unsigned i = 4; i < TL_W; ++i) fprintf(stderr, " %8.2f", r[i] ? (double)(r[i] - t0) / 1e3 : -1.0);
    fprintf(stderr, "\n");
  }
  if (last_attn_exit) fprintf(stderr, "[timeline] latest exit of any decode-attention CTA: %9.2f\n", (double)(last_attn_exit - t0) / 1e3);
}

// last <= max_records timeline records as a JSON array of [name, entry_us, wait_us, exit_us] (times relative to the first)
int Engine::timeline_json(char* out, int64_t cap, int max_records) {
  if (!tl_buf) PTTS_FAIL(PTTS_ESTATE, "timeline not enabled (PTTS_TIMELINE=1 before engine creation)");
  std::vector<unsigned long long> b(2 + TL_W * TL_MAX);
  PTTS_CUDA_CHECK(cudaDeviceSynchronize());
  PTTS_CUDA_CHECK(cudaMemcpy(b.data(), tl_buf, b.size() * 8, cudaMemcpyDeviceToHost));
  const unsigned long long total = b[0];
  const unsigned n = (unsigned)std::min<unsigned long long>(std::min<unsigned long long>(total, TL_MAX), (unsigned long long)max_records);
  std::string js = "[";
  unsigned long long t0 = 0;
  for (unsigned k = 0; k < n; ++k) {
    const unsigned long long* r = &b[2 + TL_W * (unsigned)((total - n + k) % TL_MAX)];
    if (k == 0) t0 = r[1];
    char name[64], rec[192];
    if (r[0] >> 28) snprintf(name, sizeof name, "gemm N=%llu kb=%llu", (r[0] >> 8) & 0xFFFFF, r[0] & 0xFF);
    else snprintf(name, sizeof name, "%s", r[0] == 2 ? "attn_decode" : r[0] == 3 ? "ln_rows" : r[0] == 4 ? "final_norm_eos" : r[0] == 5 ? "flow_embed" : r[0] == 6 ? "head_cluster" : "?");
    snprintf(rec, sizeof rec, "%s[\"%s\",%.3f,%.3f,%.3f]", k ? "," : "", name, (double)(r[1] - t0) / 1e3,
             r[2] ? (double)(r[2] - t0) / 1e3 : -1.0, r[3] ? (double)(r[3] - t0) / 1e3 : -1.0);
    js += rec;
  }
  js += "]";
  if ((int64_t)js.size() + 1 > cap) PTTS_FAIL(PTTS_EINVAL, "timeline_json: buffer of %lld bytes too small", (long long)cap);
  memcpy(out, js.c_str(), js.size() + 1);
  return PTTS_OK;
}

Engine::~Engine() {
  if (tl_buf) { tl_dump(tl_buf); tl_enable_kernels(nullptr); tl_enable_gemm_tc(nullptr); tl_enable_attention(nullptr); tl_enable_head_tc(nullptr); tl_enable_head_ss(nullptr); }
  for (auto& kv : graphs)
    if (kv.second.exec) cudaGraphExecDestroy(kv.second.exec);
  for (void* p : allocs) cudaFree(p);
  if (egress_buf) cudaFree(egress_buf);
  if (enc_scratch) cudaFree(enc_scratch);
  if (h_pinned) cudaFreeHost(h_pinned);
  if (h_mega_diag) cudaFreeHost(h_mega_diag);
  if (ev_fork) cudaEventDestroy(ev_fork);
  if (ev_join) cudaEventDestroy(ev_join);
  for (int k = 0; k < 2; ++k) {
    for (int j = 0; j < 8; ++j)
      if (ev_head[k][j]) cudaEventDestroy(ev_head[k][j]);
    for (cudaEvent_t e : {ev_lat[k], ev_mimi[k], ev_copied_lat[k], ev_copied_pcm[k]})
      if (e) cudaEventDestroy(e);
  }
  if (copy_stream) cudaStreamDestroy(copy_stream);
  if (ev_side_fork) cudaEventDestroy(ev_side_fork);
  if (ev_side_join) cudaEventDestroy(ev_side_join);
  if (side_stream) cudaStreamDestroy(side_stream);
  if (stream2) cudaStreamDestroy(stream2);
  if (stream) cudaStreamDestroy(stream);
}

void* Engine::act_alloc(size_t elems) {
  void* p = nullptr;
  if (cudaMalloc(&p, elems * act_size + 256) != cudaSuccess) return nullptr;
  cudaMemsetAsync(p, 0, elems * act_size + 256, stream);
  allocs.push_back(p);
  return p;
}

int Engine::init() {
  const ptts_config& c = cfg;
  if (c.d_model != c.n_heads * 64 || c.mimi_d != c.mimi_heads * 64)
    PTTS_FAIL(PTTS_EINVAL, "head_dim must be 64 (d_model=%d heads=%d)", c.d_model, c.n_heads);
  if (c.d_model > 1024 || c.mimi_d > 1024 || c.flow_dim > 512 || c.ldim > 64 || c.ldim % 16)
    PTTS_FAIL(PTTS_EINVAL, "unsupported widths");
  if (c.max_streams < 1 || c.kv_capacity < 8 || c.n_layers < 1)
    PTTS_FAIL(PTTS_EINVAL, "bad sizing (max_streams=%d kv_capacity=%d)", c.max_streams, c.kv_capacity);
  if (c.precision != PTTS_PRECISION_FP32 && c.precision != PTTS_PRECISION_BF16)
    PTTS_FAIL(PTTS_EINVAL, "unknown precision %d", c.precision);
  PTTS_CUDA_CHECK(cudaSetDevice(c.device));
  // the FlowLM chain (many short dependent kernels) outranks the Mimi branch (few wide kernels): its
  // CTAs are dispatched first whenever an SM frees up, so the wide SEANet grids do not stall it
  int prio_lo = 0, prio_hi = 0;
  PTTS_CUDA_CHECK(cudaDeviceGetStreamPriorityRange(&prio_lo, &prio_hi));
  PTTS_CUDA_CHECK(cudaStreamCreateWithPriority(&stream, cudaStreamNonBlocking, prio_hi));
  const bool mimi_hi = [] { const char* e = getenv("PTTS_MIMI_PRIO"); return e && e[0] == '1'; }();
  PTTS_CUDA_CHECK(cudaStreamCreateWithPriority(&stream2, cudaStreamNonBlocking, mimi_hi ? prio_hi : prio_lo));
  PTTS_CUDA_CHECK(cudaEventCreateWithFlags(&ev_fork, cudaEventDisableTiming));
  PTTS_CUDA_CHECK(cudaEventCreateWithFlags(&ev_join, cudaEventDisableTiming));
  if (const char* f = getenv("PTTS_PIPELINE")) use_pipeline = !(f[0] == '0');
  if (const char* f = getenv("PTTS_LOOP")) loop_fork = f[0] == 'f';
  PTTS_CUDA_CHECK(cudaDeviceGetAttribute(&n_sms, cudaDevAttrMultiProcessorCount, c.device));
  n_sms_flow = n_sms_mimi = n_sms;
  // PTTS_MIMI_GRID_SMS=k: size the grids of the Mimi branch for k SMs (no partition: it only leaves SMs idle for the FlowLM chain)
  if (const char* f = getenv("PTTS_MIMI_GRID_SMS")) { const int k = atoi(f); if (k >= 8 && k <= n_sms) n_sms_mimi = k; }
  {  // PTTS_FLOW_SMS=k: FlowLM chain on k SMs, Mimi branch on the others (green contexts); 0 / unset: shared SMs
    const char* e = getenv("PTTS_FLOW_SMS");
    const int want = e ? atoi(e) : 0;
    if (want >= 8 && want <= n_sms - 8) {
      cudaStream_t a = nullptr, b = nullptr;
      int na = 0, nb = 0;
      if (make_sm_partition(c.device, want, prio_hi, prio_lo, &a, &b, &na, &nb)) {
        cudaStreamDestroy(stream); cudaStreamDestroy(stream2);
        stream = a; stream2 = b; n_sms_flow = na; n_sms_mimi = nb; n_sms = na; partitioned = true;
      }
    }
  }
  // developer timing probes that produce WRONG audio (skipped kernels, no GELU, one branch of the loop only): never
  // silent - every engine created under them says so on stderr
  for (const char* probe : {"PTTS_SKIP_TAGS", "PTTS_DEBUG_NOGELU", "PTTS_PROFILE_PART"})
    if (const char* v = getenv(probe))
      if (v[0] && !(probe[5] == 'P' && v[0] == '0'))
        fprintf(stderr, "[pocket_tts_b200] WARNING: %s=%s is set - developer timing probe, outputs of this engine are NOT valid audio\n", probe, v);
  if (const char* f = getenv("PTTS_FORCE_SIMT")) use_tc = !(f[0] == '1');
  if (const char* f = getenv("PTTS_MIMI_GROUP")) mimi_gmax = std::max(1, std::min(8, atoi(f)));
  PTTS_CUDA_CHECK(cudaStreamCreateWithPriority(&copy_stream, cudaStreamNonBlocking, prio_lo));
  PTTS_CUDA_CHECK(cudaStreamCreateWithPriority(&side_stream, cudaStreamNonBlocking, prio_hi));
  PTTS_CUDA_CHECK(cudaEventCreateWithFlags(&ev_side_fork, cudaEventDisableTiming));
  PTTS_CUDA_CHECK(cudaEventCreateWithFlags(&ev_side_join, cudaEventDisableTiming));
  if (const char* f = getenv("PTTS_SIDE")) use_side_branch = !(f[0] == '0');
  if (const char* f = getenv("PTTS_TIMELINE")) {
    if (f[0] == '1') {
      tl_buf = dmalloc<unsigned long long>(allocs, 2 + TL_W * TL_MAX, stream);
      if (tl_buf) { tl_enable_kernels(tl_buf); tl_enable_gemm_tc(tl_buf); tl_enable_attention(tl_buf); tl_enable_head_tc(tl_buf); tl_enable_head_ss(tl_buf); }
    }
  }
  if (const char* f = getenv("PTTS_KV_SIDE")) use_kv_side_prefetch = f[0] == '1';
  if (const char* f = getenv("PTTS_LNFOLD")) use_ln_fold = !(f[0] == '0');
  if (const char* f = getenv("PTTS_HEAD_GATE")) use_head_gate = f[0] == '1';
  if (const char* f = getenv("PTTS_HEAD_GATE_MAX")) head_gate_max = atoi(f);
  if (const char* f = getenv("PTTS_SEANET_FUSED")) use_fused_stage2 = !(f[0] == '0');
  if (const char* f = getenv("PTTS_FUSED_SMS")) fused_sms = atoi(f);
  if (const char* f = getenv("PTTS_RESBLOCK_FUSED")) use_fused_resblock = !(f[0] == '0');
  if (const char* f = getenv("PTTS_HEAD")) head_mode = (f[0] == '0' || f[0] == '2' || f[0] == '3') ? f[0] - '0' : 3;
  for (int k = 0; k < 2; ++k) {
    for (cudaEvent_t* e : {&ev_lat[k], &ev_mimi[k], &ev_copied_lat[k], &ev_copied_pcm[k]})
      PTTS_CUDA_CHECK(cudaEventCreateWithFlags(e, cudaEventDisableTiming));
    for (int j = 0; j < 8; ++j) PTTS_CUDA_CHECK(cudaEventCreateWithFlags(&ev_head[k][j], cudaEventDisableTiming));
  }
  act_type = c.precision == PTTS_PRECISION_BF16 ? 1 : 0;
  act_size = act_type ? 2 : 4;
  const int S = c.max_streams;
  const int U = c.upsample_stride;

  // ---- per-slot state
  kv_flow = KvGeom{nullptr, S, c.n_heads, c.kv_capacity, 64, act_type, 0, 1};
  if (const char* f = getenv("PTTS_ATTN")) kv_flow.prefetch = f[0] >= '0' && f[0] <= '2' ? f[0] - '0' : 1;   // 0 per-lane loads, 1 + L2 prefetch before the wait, 2 bulk-copy kernel
  const size_t kvf = (size_t)c.n_layers * S * 2 * c.n_heads * c.kv_capacity * 64;
  if (c.kv_fp8) {
    if (act_type != 1) PTTS_FAIL(PTTS_EINVAL, "kv_fp8 needs PTTS_PRECISION_BF16");
    kv_flow.type = 2;
    kv_flow.base = dmalloc<uint8_t>(allocs, kvf, stream);
    kv_flow.scale = dmalloc<float>(allocs, kvf / 64, stream);
    kv_flow.prefetch = 1;
    if (!kv_flow.scale) PTTS_FAIL(PTTS_ECUDA, "out of device memory for stream state");
  } else {
    kv_flow.base = act_alloc(kvf);
  }
  int ring = 256;
  while (ring < c.mimi_context + U) ring *= 2;
  kv_mimi = KvGeom{nullptr, S, c.mimi_heads, ring, 64, act_type, 1, 0};
  kv_mimi.base = act_alloc((size_t)c.mimi_layers * S * 2 * c.mimi_heads * ring * 64);
  offsets = dmalloc<int>(allocs, S, stream);
  mimi_offsets = dmalloc<int>(allocs, S, stream);
  prev_latent = dmalloc<float>(allocs, (size_t)S * c.ldim, stream);
  e_prev = dmalloc<float>(allocs, (size_t)S * c.mimi_d, stream);
  if (!kv_flow.base || !kv_mimi.base || !offsets || !mimi_offsets || !prev_latent || !e_prev)
    PTTS_FAIL(PTTS_ECUDA, "out of device memory for stream state");
  host_offsets.assign(S, 0);
  host_mimi_offsets.assign(S, 0);
  host_pfx_slot.assign(S, -1);
  host_pfx_len.assign(S, 0);
  pfx_refs.assign(S, 0);
  d_pfx_slot = dmalloc<int>(allocs, S, stream);
  d_pfx_len = dmalloc<int>(allocs, S, stream);
  if (!d_pfx_slot || !d_pfx_len) PTTS_FAIL(PTTS_ECUDA, "out of device memory for stream state");
  PTTS_CUDA_CHECK(cudaMemsetAsync(d_pfx_slot, 0xff, S * sizeof(int), stream));
  slot_used.assign(S, 0);

  // ---- per-batch buffers
  max_rows = std::max(std::max(S, c.max_prefill_rows), 1);
  const int D = c.d_model, F = c.flow_dim, Md = c.mimi_d;
  d_slot_ids = dmalloc<int>(allocs, S, stream);
  d_row_start = dmalloc<int>(allocs, S, stream);
  d_tokens = dmalloc<int>(allocs, max_rows, stream);
  d_lens = dmalloc<int>(allocs, S, stream);
  d_mimi_row_start = dmalloc<int>(allocs, S, stream);
  d_mimi_lens = dmalloc<int>(allocs, S, stream);
  h = dmalloc<float>(allocs, (size_t)max_rows * D, stream);
  qkv = dmalloc<float>(allocs, (size_t)max_rows * 3 * D, stream);
  a_ln = act_alloc((size_t)max_rows * D);
  attn_o = act_alloc((size_t)max_rows * D);
  if (act_type == 1) {
    head_xchg = dmalloc<bf16>(allocs, (size_t)((S + 63) / 64) * 2 * 8 * 64 * 64, stream);
    if (const char* f = getenv("PTTS_HEAD_XCHG")) head_l2_exchange = !(f[0] == 'd');
    h16 = dmalloc<bf16>(allocs, (size_t)S * D, stream);
    ln_part = dmalloc<float2>(allocs, (size_t)S * (D / 16), stream);
    if (!h16 || !ln_part) PTTS_FAIL(PTTS_ECUDA, "out of device memory for activations");
  }
  mid = act_alloc((size_t)max_rows * c.ffn);
  eos_logit = dmalloc<float>(allocs, S, stream);
  noise_buf = dmalloc<float>(allocs, (size_t)S * c.ldim, stream);
  lat_in = dmalloc<float>(allocs, (size_t)S * c.ldim, stream);
  x_lat = dmalloc<float>(allocs, (size_t)S * c.ldim, stream);
  const int GM = mimi_gmax;
  for (int k = 0; k < 2; ++k) {
    lat_ring[k] = dmalloc<float>(allocs, (size_t)GM * S * c.ldim, stream);
    eos_grp[k] = dmalloc<float>(allocs, (size_t)GM * S, stream);
  }
  noise_grp = dmalloc<float>(allocs, (size_t)GM * S * c.ldim, stream);
  c_cond = act_alloc((size_t)S * D);
  cemb = dmalloc<float>(allocs, (size_t)S * F, stream);
  sy = act_alloc((size_t)S * F);
  mod = dmalloc<float>(allocs, (size_t)S * (c.flow_depth * 3 + 2) * F, stream);
  u = dmalloc<float>(allocs, (size_t)S * F, stream);
  g = act_alloc((size_t)S * F);
  g1 = act_alloc((size_t)S * F);
  xm = dmalloc<float>(allocs, (size_t)S * GM * U * Md, stream);
  mqkv = dmalloc<float>(allocs, (size_t)S * GM * U * 3 * Md, stream);
  m_ln = act_alloc((size_t)S * GM * U * Md);
  m_attn = act_alloc((size_t)S * GM * U * Md);
  m_mid = act_alloc((size_t)S * GM * U * c.mimi_ffn);
  if (!h || !qkv || !a_ln || !attn_o || !mid || !mod || !xm || !mqkv || !m_ln || !m_attn || !m_mid)
    PTTS_FAIL(PTTS_ECUDA, "out of device memory for activations");

  // ---- conv chain buffers + history table
  int C = c.n_filters * 8;  // 512
  if (C != Md) PTTS_FAIL(PTTS_EINVAL, "seanet width %d != mimi_d %d", C, Md);
  int T = U;
  hist_tab.count = 0;
  auto add_hist = [&](void* buf, int hist, int Tn, int Cn) {
    HistDesc& hd = hist_tab.d[hist_tab.count];
    hd.buf = buf; hd.hist = hist; hd.T = Tn; hd.C = Cn;
    hd.state = act_alloc((size_t)S * hist * Cn);
    hist_state[hist_tab.count] = hd.state;
    hist_tab.count++;
  };
  // conv chain buffers hold up to GM frames per stream; hist_tab records the single-frame geometry
  b_conv0 = act_alloc((size_t)S * (c.kernel - 1 + GM * T) * C);
  add_hist(b_conv0, c.kernel - 1, T, C);
  for (int i = 0; i < 3; ++i) {
    const int r = c.ratios[i], T2 = T * r, C2 = C / 2, hc = C2 / 2;
    b_tr[i] = act_alloc((size_t)S * (1 + GM * T) * C);
    add_hist(b_tr[i], 1, T, C);
    b_raw[i] = act_alloc((size_t)S * GM * T2 * C2);
    b_res[i] = act_alloc((size_t)S * (c.res_kernel - 1 + GM * T2) * C2);
    add_hist(b_res[i], c.res_kernel - 1, T2, C2);
    b_hid[i] = act_alloc((size_t)S * GM * T2 * hid_width(hc));
    T = T2; C = C2;
  }
  b_last = act_alloc((size_t)S * (c.last_kernel - 1 + GM * T) * C);
  add_hist(b_last, c.last_kernel - 1, T, C);
  pcm = dmalloc<float>(allocs, (size_t)S * GM * T, stream);
  pcm_ring[0] = dmalloc<float>(allocs, (size_t)S * GM * T, stream);
  pcm_ring[1] = dmalloc<float>(allocs, (size_t)S * GM * T, stream);
  if (!b_last || !pcm) PTTS_FAIL(PTTS_ECUDA, "out of device memory for conv buffers");
  PTTS_CUDA_CHECK(cudaMallocHost(&h_pinned, (size_t)S * 4 * sizeof(int) + 64));
  { int rc = mega_init(); if (rc != PTTS_OK) return rc; }
  PTTS_CUDA_CHECK(cudaStreamSynchronize(stream));
  return PTTS_OK;
}

// -------------------------------------------------------------------------------------
// weights
// -------------------------------------------------------------------------------------
int Engine::load_tensor(const char* name, const float* data, const int64_t* shape, int ndim) {
  if (ready) PTTS_FAIL(PTTS_ESTATE, "load_tensor after finalize");
  if (!name || !data || ndim < 0 || ndim > 4) PTTS_FAIL(PTTS_EINVAL, "bad load_tensor arguments");
  std::string nm(name);
  HostTensor t;
  t.shape.assign(shape, shape + ndim);
  const int64_t n = t.numel();
  if (n <= 0 || n > (1LL << 31)) PTTS_FAIL(PTTS_EINVAL, "bad shape for %s", name);
  t.data.assign(data, data + n);
  host_w[nm] = std::move(t);
  return PTTS_OK;
}

float* Engine::upload(const std::vector<float>& host) {
  float* p = dmalloc<float>(allocs, host.size(), stream);
  if (!p) return nullptr;
  if (cudaMemcpyAsync(p, host.data(), host.size() * sizeof(float), cudaMemcpyHostToDevice, stream) != cudaSuccess)
    return nullptr;
  cudaStreamSynchronize(stream);  // host vector may be a temporary
  return p;
}

float* Engine::dev_f32(const std::string& name, const std::vector<int64_t>& shape) {
  auto it = host_w.find(name);
  if (it == host_w.end()) { set_error("missing tensor " + name); return nullptr; }
  if (it->second.shape != shape) {
    std::string s = "shape mismatch for " + name + ": got [";
    for (auto v : it->second.shape) s += std::to_string(v) + ",";
    s += "] expected [";
    for (auto v : shape) s += std::to_string(v) + ",";
    set_error(s + "]");
    return nullptr;
  }
  float* p = upload(it->second.data);
  if (!p) set_error("device upload failed for " + name);
  return p;
}

int Engine::make_mat_from(const std::vector<float>& host, int N, int K, Mat* m, bool want_b16) {
  if ((int64_t)host.size() != (int64_t)N * K) PTTS_FAIL(PTTS_EINVAL, "matrix size mismatch");
  m->N = N; m->K = K;
  m->f32 = upload(host);
  if (!m->f32) PTTS_FAIL(PTTS_ECUDA, "upload failed");
  if (act_type == 1 && want_b16) {
    m->b16 = dmalloc<bf16>(allocs, host.size(), stream);
    if (!m->b16) PTTS_FAIL(PTTS_ECUDA, "alloc failed");
    PTTS_CUDA_CHECK(launch_f32_to_bf16(m->f32, m->b16, (long long)host.size(), stream));
  }
  return PTTS_OK;
}

int Engine::make_mat(const std::string& name, int N, int K, Mat* m, bool want_b16) {
  auto it = host_w.find(name);
  if (it == host_w.end()) PTTS_FAIL(PTTS_EINVAL, "missing tensor %s", name.c_str());
  if (it->second.numel() != (int64_t)N * K || it->second.shape.empty() || it->second.shape[0] != N)
    PTTS_FAIL(PTTS_EINVAL, "shape mismatch for %s (want [%d,%d])", name.c_str(), N, K);
  return make_mat_from(it->second.data, N, K, m, want_b16);
}

}  // namespace ptts
