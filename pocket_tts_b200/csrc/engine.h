// Engine: weights, per-slot state (FlowLM KV slabs, Mimi ring KV, conv history), per-batch
// activation buffers and the kernel sequences of the frame step.  Host-side C++ only sees
// this header; the C ABI (c_api.cu) is a thin wrapper.
#pragma once
#include <cuda_runtime.h>

#include <cstdio>
#include <functional>
#include <map>
#include <string>
#include <vector>

#include "../../include/pocket_tts_b200.h"
#include "flow_mega.h"
#include "gemm.cuh"
#include "kernels.cuh"

namespace ptts {

void set_error(const std::string& msg);
void set_error_cuda(cudaError_t e, const char* expr, const char* file, int line);
const char* get_error();

struct HostTensor {
  std::vector<float> data;
  std::vector<int64_t> shape;
  int64_t numel() const { int64_t n = 1; for (auto s : shape) n *= s; return n; }
};

// a GEMM weight [N, K] row-major in device memory, fp32 and (bf16 mode) bf16
struct Mat {
  float* f32 = nullptr;
  bf16* b16 = nullptr;
  int N = 0, K = 0;
};

struct LayerW {
  float *n1w, *n1b, *n2w, *n2b;
  Mat in_proj, out_proj, lin1, lin2;
  float *ls1 = nullptr, *ls2 = nullptr;  // Mimi LayerScale
  // LayerNorm folded into the decode GEMMs (bf16 mode, FlowLM): W diag(gamma) in bf16, c1 = its row sums, c2 = W beta
  Mat in_proj_ln, lin1_ln;
  float *qkv_c1 = nullptr, *qkv_c2 = nullptr, *ffn_c1 = nullptr, *ffn_c2 = nullptr;
};

struct FlowBlockW {
  float *lnw, *lnb, *b0, *b2;
  Mat mlp0, mlp2;
};

struct ConvStageW {
  Mat tr;  float* tr_b;      // [S*Cout, 2Cin], bias expanded to [S*Cout]
  Mat c3;  float* c3_b;      // [h, 3*C2]
  Mat c1;  float* c1_b;      // [C2, h]
};

// Mimi encoder side (voice-prompt encoding, engine_encoder.cu); all fp32
struct EncStageW {
  Mat c3;   float* c3_b = nullptr;     // ResBlock conv K3 [C/2, 3C]
  Mat c1;   float* c1_b = nullptr;     // ResBlock conv K1 [C, C/2]
  Mat down; float* down_b = nullptr;   // strided conv K = 2r, S = r: [2C, 2r*C]
  int ratio = 0;
};
struct EncoderW {
  bool loaded = false;
  float *w0 = nullptr, *b0 = nullptr;  // first conv [C0, 1, K] as stored
  EncStageW st[3];
  Mat last; float* last_b = nullptr;   // final conv K3 -> mimi_d
  std::vector<LayerW> layers;          // encoder transformer
  Mat down;                            // ConvDownsample1d [ldim, 2U*mimi_d]
  Mat speaker;                         // flow_lm.speaker_proj_weight [d_model, ldim]
};

struct MegaPlan {            // phase program of the persistent FlowLM step kernel for one lsd_steps value
  CUtensorMap* d_maps = nullptr;
  MegaPhase* d_phases = nullptr;
  int n_phases = 0;
};
const unsigned* mega_diag_words();
void tl_enable_kernels(unsigned long long*);
void tl_enable_gemm_tc(unsigned long long*);
void tl_enable_attention(unsigned long long*);
void tl_enable_head_tc(unsigned long long*);
void tl_enable_head_ss(unsigned long long*);
// static_scratch.cu: PCM egress conversion staged through a file-scope device array (no allocation)
int pcm_to_int16_static(cudaStream_t stream, const float* pcm_any, short* out_any, long long n, long long* launches);
bool make_sm_partition(int device, int flow_sms, int prio_flow, int prio_mimi, cudaStream_t* s_flow, cudaStream_t* s_mimi,
                       int* n_flow, int* n_mimi);

struct GraphKey {
  int n, lsd, audio;
  int mode = 0;   // 0: sequential frame; 8 + parity*4 + do_flow*2 + do_mimi: pipelined frame
  int shared = 0; // 1: captured with the shared-prefix attention kernels (KvGeom pointers are baked into the nodes)
  bool operator<(const GraphKey& o) const {
    if (shared != o.shared) return shared < o.shared;
    if (n != o.n) return n < o.n;
    if (lsd != o.lsd) return lsd < o.lsd;
    if (audio != o.audio) return audio < o.audio;
    return mode < o.mode;
  }
};
struct GraphEntry {
  cudaGraphExec_t exec = nullptr;
  long long kernels = 0;
};

constexpr int MIMI_SEGS = 5;   // segments of a Mimi pass that can be captured separately (Engine::mimi_sequence)

class Engine {
 public:
  explicit Engine(const ptts_config& c) : cfg(c) {}
  ~Engine();
  int init();
  int load_tensor(const char* name, const float* data, const int64_t* shape, int ndim);
  int finalize();

  int stream_open(int* slot);
  int stream_close(int slot);
  int stream_reset_decoder(int slot);
  int stream_fork(int src, int dst);
  int stream_fork_shared(int src, int dst);
  int import_kv(int slot, int layer, const float* cache_any, int P);
  int export_kv(int slot, int layer, float* cache_any, int P);
  int set_offset(int slot, int off);
  int get_offset(int slot, int* off);
  int set_prev_latent(int slot, const float* latent_any);

  int prefill(int n, const int* slots, const int* lens, const float* embeds_any, const int* tokens = nullptr);
  int step(int n, const int* slots, const float* noise_any, int lsd, int audio, float* out_lat,
           float* out_eos, float* out_pcm);
  int mimi_decode(int n, const int* slots, const float* latents_any, float* out_pcm);
  int mimi_decode_frames(int n, const int* slots, int n_frames, const float* latents_any, float* out_pcm_any);
  int run_frames(int n, const int* slots, int n_frames, const float* noise_any, int lsd,
                 float* out_pcm_any, float* out_lat_any, float* out_eos_any = nullptr, bool sync = false);
  int debug_read(const char* what, float* out_any, int64_t max_elems, int64_t* n_out);
  int debug_gemm(int backend, int M, int N, int K, int ld, int rows_per_batch, int valid_rows,
                 const float* a_host, long long a_elems, const float* w_host, float* out_host);
  int pcm_to_int16(const float* pcm_any, short* out_any, long long n);
  int encode_audio(const float* audio_any, long long n_samples, float* out_cond_any, int max_frames, int* n_frames);
  int timeline_json(char* out, int64_t cap, int max_records);
  int profile_begin();
  int profile_end(char* json_out, int64_t cap);

  ptts_config cfg;
  cudaStream_t stream = nullptr;
  cudaStream_t stream2 = nullptr;          // Mimi branch of the pipelined frame
  cudaEvent_t ev_fork = nullptr, ev_join = nullptr;
  bool use_pipeline = true;                // PTTS_PIPELINE=0 disables
  bool loop_fork = false;                  // PTTS_LOOP=fork: one two-branch graph per group instead of per-frame FlowLM graphs + per-group Mimi graphs
  long long launches = 0;
  bool use_graphs = true;
  bool use_tc = true;   // bf16 mode: tcgen05 GEMMs (PTTS_FORCE_SIMT=1 disables)
  int n_sms = 148;                         // SMs the kernels of the current branch may assume (swapped with the stream)
  int n_sms_flow = 148, n_sms_mimi = 148;  // per branch; differ when the SMs are partitioned (PTTS_FLOW_SMS)
  bool partitioned = false;

 private:
  // ---- helpers
  int check_slots(int n, const int* slots, bool need_decode_room);
  int upload_slots(int n, const int* slots);
  float* dev_f32(const std::string& name, const std::vector<int64_t>& shape);
  int make_mat(const std::string& name, int N, int K, Mat* m, bool want_b16 = true);
  int make_mat_from(const std::vector<float>& host, int N, int K, Mat* m, bool want_b16 = true);
  float* upload(const std::vector<float>& host);
  int gemm(const char* tag, const void* A, bool a_is_act, RowMap amap, const Mat& W, int M, GemmEpilogue ep,
           bool out_is_act, const KvPrefetchJob* kv_prefetch = nullptr);
  KvPrefetchJob flow_kv_prefetch_job(int layer, int n) const;
  void prof_begin(const char* tag);
  void prof_end();
  int load_encoder();
  int make_ln_folded(const std::string& wname, const std::string& gname, const std::string& bname, int N, int K, Mat* m,
                     float** c1, float** c2);
  int flow_layer_decode_folded(int l, int n);
  int transformer_layer_common(const LayerW& L, float* h, int rows, int d, int ffn, void* a,
                               float* qkv, void* ao, void* mid, bool mimi, int li, int n);
  int flow_sequence(int n, int lsd, const float* noise = nullptr, float* eos_out = nullptr);
  // persistent single-launch form of flow_sequence (bf16 mode, <= 128 streams): engine_mega.cu
  int mega_init();
  bool mega_eligible(int n) const;
  int mega_plan(int lsd, MegaPlan** out);
  int flow_sequence_mega(int n, int lsd, const float* noise, float* eos_out);
  int mimi_sequence(int n, const float* latents, int frames = 1, float* pcm_out = nullptr, int seg_lo = 0, int seg_hi = MIMI_SEGS - 1);
  int frame(int n, int lsd, int audio);  // eager or graph
  // FlowLM + head of the current frame on `stream` in parallel with the Mimi decode of the previous
  // frame's latents on `stream2` (two branches of one graph), like the reference's decoder thread
  // (tts_model.py:651-658): the Mimi chain never feeds back into the FlowLM chain.
  int launch_captured(const GraphKey& key, cudaStream_t on, const std::function<int()>& body);
  int flow_graph(int n, int lsd, int par, int j);
  int mimi_graph(int n, int g, int par);
  int mimi_graph_gated(int n, int g, int par, int gate_par);
  int group_graph(int n, int lsd, int parity, int g_flow, int g_mimi);
  void* act_alloc(size_t elems);
  // ResBlock hidden width as stored (bf16 mode pads to a multiple of 64 channels)
  int hid_width(int hc) const { return act_type == 1 ? ((hc + 63) / 64) * 64 : hc; }

  bool use_mega = false;                   // PTTS_MEGA=1 selects the persistent step kernel (experimental: see DESIGN.md)
  int mega_ctas = 128, mega_stages = 6, mega_rows_pad = 64;
  unsigned *d_mega_sync = nullptr, *h_mega_diag = nullptr, *d_mega_diag = nullptr;
  float* mega_part = nullptr;
  unsigned long long* d_mega_trace = nullptr;
  std::map<int, MegaPlan> mega_plans;

  EncoderW enc;
  float* enc_scratch = nullptr;              // encode_audio working set, grown on demand
  long long enc_scratch_elems = 0;
  bool ready = false;
  int act_type = 0;  // 0 fp32, 1 bf16
  size_t act_size = 4;
  std::map<std::string, HostTensor> host_w;
  std::vector<void*> allocs;

  // weights
  float *bos_emb, *w_in, *out_norm_w, *out_norm_b, *eos_w, *eos_b, *emb_std, *emb_mean;
  float *rope_freqs, *tcomb;
  float* text_embed = nullptr;               // optional LUT conditioner table [text_embed_rows, d_model] (ptts_prefill_tokens)
  int text_embed_rows = 0;
  int* d_tokens = nullptr;                   // [max_rows]
  int tcomb_rows = 0;
  std::vector<LayerW> flow_layers, mimi_layers;
  Mat cond_embed, input_proj, mod_all, final_lin;
  float *cond_b, *inproj_b, *mod_b, *final_b;
  float* cond_b_t0 = nullptr;   // cond_embed bias + (te(s)+te(t))/2 of LSD step 0: lsd_steps == 1 folds SiLU into the GEMM
  std::vector<FlowBlockW> flow_blocks;
  float *wq, *wup;
  Mat conv0; float* conv0_b;
  ConvStageW stages[3];
  float *last_w, *last_b;

  // per-slot state
  KvGeom kv_flow{}, kv_mimi{};
  int *offsets = nullptr, *mimi_offsets = nullptr;  // device [slots]
  float* prev_latent = nullptr;                      // device [slots, ldim]
  float* e_prev = nullptr;                           // device [slots, mimi_d]
  void* hist_state[PTTS_MAX_HIST] = {};
  std::vector<int> host_offsets, host_mimi_offsets;
  // shared voice prefixes (stream_fork_shared): per slot the owner of positions [0, pfx_len) or -1, and how many live
  // slots read a slot's prefix; mirrored on the device for the attention kernels
  std::vector<int> host_pfx_slot, host_pfx_len, pfx_refs;
  int *d_pfx_slot = nullptr, *d_pfx_len = nullptr;
  bool batch_shared = false;                 // some slot of the current batch (cur_slots) reads a shared prefix
  KvGeom flow_kv() const { KvGeom g = kv_flow; if (batch_shared) { g.pfx_slot = d_pfx_slot; g.pfx_len = d_pfx_len; } return g; }
  int unshare(int slot);
  std::vector<char> slot_used;

  // per-batch buffers
  int max_rows = 0;
  int *d_slot_ids, *d_row_start, *d_lens, *d_mimi_row_start, *d_mimi_lens;
  float *h, *qkv, *eos_logit, *noise_buf, *x_lat, *cemb, *mod, *u, *pcm, *lat_in;
  void *a_ln, *attn_o, *mid, *c_cond, *sy, *g, *g1;
  unsigned long long* tl_buf = nullptr;      // developer timeline ring (PTTS_TIMELINE=1), dumped by the destructor
  bf16* head_xchg = nullptr;                 // flow-head cluster kernel: operand exchange scratch (flow_head_tc.cu)
  bool head_l2_exchange = true;              // PTTS_HEAD_XCHG=dsmem: exchange through DSMEM bulk copies instead
  bf16* h16 = nullptr;                       // bf16 copy of the residual stream (operand of the LayerNorm-folded GEMMs)
  float2* ln_part = nullptr;                 // [rows][d_model / 16] (sum, M2) partials of the residual rows
  float *xm, *mqkv;
  void *m_ln, *m_attn, *m_mid;
  // conv chain buffers
  void* b_conv0;                  // [n][6+16][512]
  void* b_tr[3];                  // [n][1+T][C]
  void* b_raw[3];                 // [n][T2][C2]
  void* b_res[3];                 // [n][2+T2][C2]
  void* b_hid[3];                 // [n][T2][h]
  void* b_last;                   // [n][2+1920][64]
  HistTable hist_tab{};
  int frame_T0 = 16;

  // per-kernel CUDA-event profiling (eager mode only)
  struct ProfRec { const char* tag; cudaEvent_t a, b; };
  bool prof_on = false;
  std::vector<ProfRec> prof_recs;
  std::vector<cudaEvent_t> prof_pool;

  std::map<GraphKey, GraphEntry> graphs;
  std::vector<int> cur_slots;
  void* egress_buf = nullptr; size_t egress_cap = 0;   // staging of ptts_pcm_to_int16 for host pointers
  int prefill_max_len = 0;                   // longest chunk of the current prefill batch (selects the two-pass attention)
  float* h_pinned = nullptr;
  float* lat_ring[2] = {nullptr, nullptr};   // latents [mimi_gmax][S][ldim] handed from the FlowLM branch to the Mimi branch
  float* pcm_ring[2] = {nullptr, nullptr};   // PCM [S][mimi_gmax * 1920] of the group decoded by the Mimi branch
  float* noise_grp = nullptr;                // [mimi_gmax][S][ldim]
  float* eos_grp[2] = {nullptr, nullptr};    // [mimi_gmax][S]
  int mimi_gmax = 4;                         // frames per Mimi pass in the pipelined loop (PTTS_MIMI_GROUP)
  cudaStream_t side_stream = nullptr;        // side branch of the FlowLM graph (input_proj of the first Euler step)
  cudaEvent_t ev_side_fork = nullptr, ev_side_join = nullptr;
  bool use_side_branch = true;               // PTTS_SIDE=0 disables
  bool use_kv_side_prefetch = false;         // PTTS_KV_SIDE=1: the QKV GEMM's idle epilogue warps prefetch the layer's K/V rows for the attention kernel (measured: attention 13.2 -> 12.8 us but the GEMM 8.5 -> 10.6 us on the chain, so off)
  bool use_fused_stage2 = true;              // PTTS_SEANET_FUSED=0: last SEANet stage as separate GEMM launches + last_conv (the fp32 mode always does)
  int fused_sms = 0;                         // PTTS_FUSED_SMS: 0 persistent, one CTA per SM (default: 112 vs 140 us per 4-frame pass), k > 0 persistent on k SMs, -1 one tile per CTA
  FusedStage2Consts fused2_consts{};
  bool use_fused_resblock = true;            // PTTS_RESBLOCK_FUSED=0: ResBlocks of stages 0 / 1 as two GEMM launches each
  std::vector<float> res_b3_host[3], res_b1_host[3];   // host copies of the ResBlock biases (kernel parameters of the fused form)
  bool use_ln_fold = true;                   // PTTS_LNFOLD=0: separate LayerNorm launches in the FlowLM decode layers
  int head_mode = 3;                         // PTTS_HEAD: 0 flow head as separate launches (A/B reference), 2 feature-split 8-CTA cluster kernel on tcgen05 (flow_head_tc.cu), 3 stream-split kernel without cross-CTA exchange (flow_head_ss.cu, default); the first cluster form on mma.sync (1) was retired in round 2
  cudaStream_t copy_stream = nullptr;        // result copies of run_frames ride beside the next group's graph
  cudaEvent_t ev_lat[2] = {}, ev_mimi[2] = {}, ev_copied_lat[2] = {}, ev_copied_pcm[2] = {};
  // Head-gated Mimi passes: ev_head[par][j] is recorded inside the FlowLM graph of frame slot j of the
  // group with parity par when its flow head starts (8 SMs busy, 140 idle for ~65 us); the SEANet segments of the
  // previous group's Mimi pass - the kernels that fill whole SMs - wait for those events instead of colliding with the
  // transformer layers of the chain.
  cudaEvent_t ev_head[2][8] = {};
  cudaEvent_t gate_record = nullptr;         // set around the capture of a FlowLM frame graph
  bool use_head_gate = true;                 // PTTS_HEAD_GATE=0: Mimi passes run whenever their latents are ready
  int head_gate_max = 256;                   // PTTS_HEAD_GATE_MAX: largest batch the gated loop is used for (measured: +5 % at 64 streams x 6 layers, +2.5 % at 256 x 24)
  bool mimi_alone = false;                   // the Mimi pass being issued has no FlowLM chain beside it (decoder-only calls)
  bool gate_active = false;                  // the running run_frames call uses the gated loop
  float* commit_out = nullptr;               // extra destination of head.commit (pipelined frames)
  std::map<std::string, std::pair<const void*, std::pair<int, long long>>> debug_bufs;  // name -> (ptr,(type,elems per row unit))
};


template <typename T>
static inline T* dmalloc(std::vector<void*>& allocs, size_t n, cudaStream_t st) {
  void* p = nullptr;
  if (cudaMalloc(&p, n * sizeof(T) + 256) != cudaSuccess) return nullptr;
  cudaMemsetAsync(p, 0, n * sizeof(T) + 256, st);
  allocs.push_back(p);
  return reinterpret_cast<T*>(p);
}

}  // namespace ptts

#define PTTS_FAIL(code, ...)                       \
  do {                                             \
    char _b[512];                                  \
    snprintf(_b, sizeof _b, __VA_ARGS__);          \
    ptts::set_error(_b);                           \
    return code;                                   \
  } while (0)
