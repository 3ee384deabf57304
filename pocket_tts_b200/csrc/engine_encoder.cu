// Engine implementation, part 5: the Mimi ENCODER side on the device — voice-prompt
// encoding for TTSModel.get_state_for_audio_prompt:
//   TTSModel._encode_audio         models/tts_model.py:379-388
//   MimiModel.encode_to_latent     models/mimi.py:96-119  (right padding to whole frames: modules/conv.py:14-33)
//   SEANetEncoder                  modules/seanet.py:44-113  (ratios reversed: 4, 5, 6; one ResBlock per stage)
//   encoder ProjectedTransformer   modules/mimi_transformer.py:129-150 (no projections: 512 == 512), causal,
//                                  context 250, LayerScale, RoPE from position 0 (model_state=None)
//   ConvDownsample1d               modules/resample.py:7-29  (K = 32, S = 16, no bias, "replicate" left padding)
//   F.linear(latents, speaker_proj_weight)                              tts_model.py:386-387
//
// Everything runs in fp32 (FFMA GEMMs, fp32 activations) whatever the engine's precision mode: it is a one-off per
// voice, its output is conditioning that every later frame depends on, and the parity gate is the fp32 one
// (rel-L2 <= 1e-4 against the reference's own _encode_audio, tests/golden/encode_*.npz).
//
// Data layout: one waveform per call, channels-last [time][channels] fp32 buffers.  Every convolution is the same
// overlapping-row GEMM as on the decoder side: a Conv1d with kernel K and stride S over a buffer with
// K - S history columns in front is a plain GEMM whose A row j = the K*C_in contiguous elements starting at j*S*C_in
// (row stride S*C_in < row length), so strided convolutions need no gather either.  The first conv (1 -> 64
// channels, K = 7) and the attention have their own small kernels.
#include <algorithm>

#include "engine.h"

namespace ptts {

static inline RowMap plain(int ld) { return RowMap{0, 1 << 30, 1 << 30, ld}; }

// ---- first conv: y[t][c] = b[c] + sum_k w[c][k] * x[t + k - (K-1)]  (zero history), raw + ELU copies
__global__ void __launch_bounds__(256)
enc_conv0_kernel(const float* __restrict__ audio, long long n_valid, long long T, const float* __restrict__ w /*[C][K]*/,
                 const float* __restrict__ b, float* __restrict__ raw /*[T][C]*/, float* __restrict__ act /*[T][C] = ELU*/,
                 int C, int K) {
  __shared__ float ws[64 * 8], bs[64];
  for (int i = threadIdx.x; i < C * K; i += blockDim.x) ws[i] = w[i];
  for (int i = threadIdx.x; i < C; i += blockDim.x) bs[i] = b[i];
  __syncthreads();
  const long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x;   // (t, c)
  if (idx >= T * C) return;
  const long long t = idx / C;
  const int c = (int)(idx - t * C);
  float acc = bs[c];
  for (int k = 0; k < K; ++k) {
    const long long s = t + k - (K - 1);
    const float x = (s >= 0 && s < n_valid) ? audio[s] : 0.0f;    // left history zeros, right padding zeros (conv.py:33)
    acc = fmaf(ws[c * K + k], x, acc);
  }
  raw[idx] = acc;
  act[idx] = elu_f(acc);
}

// rows [0, n_rows) of dst (row length C) = copies of row `src_row` (replicate padding, conv.py:101-106)
__global__ void enc_replicate_rows_kernel(float* __restrict__ dst, const float* __restrict__ src_row, int n_rows, int C) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n_rows * C) dst[i] = src_row[i % C];
}

// Causal sliding-window attention over a whole sequence held in qkv [T][3*H*64] (q | k | v, head-major inside each),
// RoPE (interleaved pairs, rope.py:28-58) applied to q and k at their own positions while staging.  One CTA per
// (block of QB queries, head): the K / V rows the block can see ([t0 - context + 1, t0 + QB)) are staged once in
// shared memory; one warp per query at a time: lanes over keys for the scores, lanes over dims for the output.
constexpr int ENC_QB = 16;
__global__ void __launch_bounds__(256)
enc_attn_kernel(const float* __restrict__ qkv, const float* __restrict__ freqs, float* __restrict__ out /*[T][H*64]*/,
                int T, int H, int context) {
  extern __shared__ float sm[];
  const int NK = context - 1 + ENC_QB;            // staged keys at most
  float* Ks = sm;                                  // [NK][65]
  float* Vs = Ks + (((size_t)NK * 65 + 3) & ~(size_t)3);   // [NK][64], 16-byte aligned
  float* Qs = Vs + (size_t)NK * 64;                // [QB][64]
  float* Ps = Qs + ENC_QB * 64;                    // [8 warps][NK]
  const int h = blockIdx.y, t0 = blockIdx.x * ENC_QB;
  const int d = H * 64, ld = 3 * d;
  const int k_lo = max(0, t0 - (context - 1)), k_hi = min(T, t0 + ENC_QB);   // staged keys [k_lo, k_hi)
  const int nk = k_hi - k_lo;
  // stage K (rotated), V: one (row, pair) per thread step
  for (int i = threadIdx.x; i < nk * 32; i += blockDim.x) {
    const int r = i >> 5, p = i & 31, pos = k_lo + r;
    const float* kr = qkv + (size_t)pos * ld + d + h * 64 + 2 * p;
    const float* vr = qkv + (size_t)pos * ld + 2 * d + h * 64 + 2 * p;
    float sn, cs;
    sincosf((float)pos * freqs[p], &sn, &cs);
    const float2 kv = *reinterpret_cast<const float2*>(kr);
    Ks[r * 65 + 2 * p] = kv.x * cs - kv.y * sn;
    Ks[r * 65 + 2 * p + 1] = kv.x * sn + kv.y * cs;
    const float2 vv = *reinterpret_cast<const float2*>(vr);
    Vs[r * 64 + 2 * p] = vv.x;
    Vs[r * 64 + 2 * p + 1] = vv.y;
  }
  for (int i = threadIdx.x; i < ENC_QB * 32; i += blockDim.x) {
    const int r = i >> 5, p = i & 31, pos = t0 + r;
    if (pos < T) {
      const float2 qv = *reinterpret_cast<const float2*>(qkv + (size_t)pos * ld + h * 64 + 2 * p);
      float sn, cs;
      sincosf((float)pos * freqs[p], &sn, &cs);
      Qs[r * 64 + 2 * p] = qv.x * cs - qv.y * sn;
      Qs[r * 64 + 2 * p + 1] = qv.x * sn + qv.y * cs;
    }
  }
  __syncthreads();
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  float* P = Ps + (size_t)warp * NK;
  for (int qi = warp; qi < ENC_QB; qi += 8) {
    const int t = t0 + qi;
    if (t >= T) break;                              // uniform per warp
    const int lo = max(0, t - (context - 1)) - k_lo, hi = t - k_lo;   // staged-row range [lo, hi] (transformer.py:22-29)
    const float* q = Qs + qi * 64;
    float mx = -INFINITY;
    for (int j = lo + lane; j <= hi; j += 32) {
      const float* kr = Ks + j * 65;
      float s = 0.0f;
#pragma unroll 16
      for (int c = 0; c < 64; ++c) s = fmaf(q[c], kr[c], s);
      s *= 0.125f;                                  // 1 / sqrt(64), SDPA default scale
      P[j] = s;
      mx = fmaxf(mx, s);
    }
    mx = warp_max(mx);
    float sum = 0.0f;
    for (int j = lo + lane; j <= hi; j += 32) {
      const float e = expf(P[j] - mx);
      P[j] = e;
      sum += e;
    }
    sum = warp_sum(sum);
    __syncwarp();
    float o0 = 0.0f, o1 = 0.0f;
    for (int j = lo; j <= hi; ++j) {
      const float p = P[j];
      const float2 v = *reinterpret_cast<const float2*>(Vs + j * 64 + 2 * lane);
      o0 = fmaf(p, v.x, o0);
      o1 = fmaf(p, v.y, o1);
    }
    const float inv = 1.0f / sum;
    *reinterpret_cast<float2*>(out + (size_t)t * d + h * 64 + 2 * lane) = make_float2(o0 * inv, o1 * inv);
    __syncwarp();
  }
}

// ---------------------------------------------------------------------------------------------------------------
// weights
// ---------------------------------------------------------------------------------------------------------------
static std::vector<float> conv_to_gemm_enc(const HostTensor& w) {   // [C_out, C_in, K] -> [C_out, K*C_in]
  const int co = (int)w.shape[0], ci = (int)w.shape[1], k = (int)w.shape[2];
  std::vector<float> g((size_t)co * k * ci);
  for (int o = 0; o < co; ++o)
    for (int c = 0; c < ci; ++c)
      for (int t = 0; t < k; ++t) g[((size_t)o * k + t) * ci + c] = w.data[((size_t)o * ci + c) * k + t];
  return g;
}

// Called from finalize(): the encoder is optional (the "without voice cloning" checkpoints do not carry it,
// tts_model.py:871-872); it is complete or absent.
int Engine::load_encoder() {
  const ptts_config& c = cfg;
  if (!host_w.count("mimi.encoder.model.0.conv.weight")) return PTTS_OK;
  auto need = [&](const std::string& n) -> const HostTensor* {
    auto it = host_w.find(n);
    if (it == host_w.end()) { set_error("missing tensor " + n + " (the Mimi encoder must be loaded whole)"); return nullptr; }
    return &it->second;
  };
  auto vec = [&](const std::string& n, int64_t len, float** dst) -> int {
    const HostTensor* t = need(n);
    if (!t) return PTTS_EINVAL;
    if (t->numel() != len) PTTS_FAIL(PTTS_EINVAL, "bad size for %s", n.c_str());
    *dst = upload(t->data);
    return *dst ? PTTS_OK : PTTS_ECUDA;
  };
  auto conv = [&](const std::string& n, int co, int ci, int k, Mat* m) -> int {
    const HostTensor* t = need(n);
    if (!t) return PTTS_EINVAL;
    if (t->shape != std::vector<int64_t>{co, ci, k}) PTTS_FAIL(PTTS_EINVAL, "bad shape for %s", n.c_str());
    return make_mat_from(conv_to_gemm_enc(*t), co, k * ci, m, false);
  };
  int rc;
  EncoderW& E = enc;
  int C = c.n_filters;
  {
    const HostTensor* t = need("mimi.encoder.model.0.conv.weight");
    if (!t) return PTTS_EINVAL;
    if (t->shape != std::vector<int64_t>{C, 1, c.kernel} || C > 64 || c.kernel > 8) PTTS_FAIL(PTTS_EINVAL, "bad encoder.model.0 shape");
    E.w0 = upload(t->data);                      // [C][K] as stored
    if ((rc = vec("mimi.encoder.model.0.conv.bias", C, &E.b0)) != PTTS_OK) return rc;
  }
  int idx = 1;
  for (int i = 0; i < 3; ++i) {
    const int r = c.ratios[2 - i], hc = C / 2;   // ratios reversed (seanet.py:63)
    const std::string pr = "mimi.encoder.model." + std::to_string(idx) + ".block.";
    const std::string pd = "mimi.encoder.model." + std::to_string(idx + 2) + ".conv.";
    if ((rc = conv(pr + "1.conv.weight", hc, C, c.res_kernel, &E.st[i].c3)) != PTTS_OK) return rc;
    if ((rc = vec(pr + "1.conv.bias", hc, &E.st[i].c3_b)) != PTTS_OK) return rc;
    if ((rc = conv(pr + "3.conv.weight", C, hc, 1, &E.st[i].c1)) != PTTS_OK) return rc;
    if ((rc = vec(pr + "3.conv.bias", C, &E.st[i].c1_b)) != PTTS_OK) return rc;
    if ((rc = conv(pd + "weight", 2 * C, C, 2 * r, &E.st[i].down)) != PTTS_OK) return rc;
    if ((rc = vec(pd + "bias", 2 * C, &E.st[i].down_b)) != PTTS_OK) return rc;
    E.st[i].ratio = r;
    C *= 2;
    idx += 3;
  }
  const int Md = c.mimi_d;
  {
    const std::string pl = "mimi.encoder.model." + std::to_string(idx + 1) + ".conv.";
    if ((rc = conv(pl + "weight", Md, C, c.last_kernel, &E.last)) != PTTS_OK) return rc;
    if ((rc = vec(pl + "bias", Md, &E.last_b)) != PTTS_OK) return rc;
  }
  E.layers.resize(c.mimi_layers);
  for (int i = 0; i < c.mimi_layers; ++i) {
    const std::string p = "mimi.encoder_transformer.transformer.layers." + std::to_string(i) + ".";
    LayerW& w = E.layers[i];
    if ((rc = vec(p + "norm1.weight", Md, &w.n1w)) != PTTS_OK || (rc = vec(p + "norm1.bias", Md, &w.n1b)) != PTTS_OK ||
        (rc = vec(p + "norm2.weight", Md, &w.n2w)) != PTTS_OK || (rc = vec(p + "norm2.bias", Md, &w.n2b)) != PTTS_OK ||
        (rc = vec(p + "layer_scale_1.scale", Md, &w.ls1)) != PTTS_OK || (rc = vec(p + "layer_scale_2.scale", Md, &w.ls2)) != PTTS_OK)
      return rc;
    if ((rc = make_mat(p + "self_attn.in_proj.weight", 3 * Md, Md, &w.in_proj, false)) != PTTS_OK) return rc;
    if ((rc = make_mat(p + "self_attn.out_proj.weight", Md, Md, &w.out_proj, false)) != PTTS_OK) return rc;
    if ((rc = make_mat(p + "linear1.weight", c.mimi_ffn, Md, &w.lin1, false)) != PTTS_OK) return rc;
    if ((rc = make_mat(p + "linear2.weight", Md, c.mimi_ffn, &w.lin2, false)) != PTTS_OK) return rc;
  }
  if ((rc = conv("mimi.downsample.conv.conv.weight", c.ldim, Md, 2 * c.upsample_stride, &E.down)) != PTTS_OK) return rc;
  if ((rc = make_mat("flow_lm.speaker_proj_weight", c.d_model, c.ldim, &E.speaker, false)) != PTTS_OK) return rc;
  E.loaded = true;
  return PTTS_OK;
}

// ---------------------------------------------------------------------------------------------------------------
// encode_audio
// ---------------------------------------------------------------------------------------------------------------
#define LAUNCH(tag, expr)       \
  do {                          \
    prof_begin(tag);            \
    PTTS_CUDA_CHECK(expr);      \
    prof_end();                 \
    ++launches;                 \
  } while (0)

int Engine::encode_audio(const float* audio_any, long long n_samples, float* out_cond_any, int max_frames, int* n_frames) {
  if (!ready) PTTS_FAIL(PTTS_ESTATE, "encode_audio before finalize");
  if (!enc.loaded) PTTS_FAIL(PTTS_ESTATE, "the Mimi encoder weights were not loaded (model without voice cloning)");
  const ptts_config& c = cfg;
  const int hop = c.upsample_stride * c.ratios[0] * c.ratios[1] * c.ratios[2];    // 1920 samples per frame
  if (!audio_any || n_samples < 1 || n_samples > PTTS_ENCODE_MAX_SAMPLES)
    PTTS_FAIL(PTTS_EINVAL, "encode_audio: n_samples %lld outside [1, %d]", n_samples, PTTS_ENCODE_MAX_SAMPLES);
  const int frames = (int)((n_samples + hop - 1) / hop);                          // conv.py:14-33, kernel = stride = frame
  if (n_frames) *n_frames = frames;
  if (!out_cond_any) return PTTS_OK;                                              // size query
  if (frames > max_frames) PTTS_FAIL(PTTS_EINVAL, "encode_audio: %d frames, room for %d", frames, max_frames);
  const long long T0 = (long long)frames * hop;
  const int C0 = c.n_filters, Md = c.mimi_d, U = c.upsample_stride;
  const int rk = c.res_kernel - 1;
  // ---- scratch (grown on demand, kept for the next call): sizes in floats
  // ping / pong hold [hist + T][C] stage inputs (T*C is the same for the three ResBlocks up to the stride), hid the
  // ResBlock hidden rows, raw the skip copy
  const long long n_act = (long long)(8 + T0) * C0 + 4096;   // history rows <= 6 at any stage, T*C shrinks after stage 0
  const long long Tt = (long long)frames * U;                                     // transformer positions
  const long long n_tr = (long long)(U + Tt) * Md;
  const long long need = 4 * n_act + n_samples + n_tr * 2 + Tt * 3 * Md + Tt * (long long)c.mimi_ffn + (long long)frames * (c.ldim + c.d_model) + 1024;
  if (need > enc_scratch_elems) {
    if (enc_scratch) { PTTS_CUDA_CHECK(cudaStreamSynchronize(stream)); cudaFree(enc_scratch); enc_scratch = nullptr; enc_scratch_elems = 0; }
    PTTS_CUDA_CHECK(cudaMalloc(&enc_scratch, (size_t)need * sizeof(float)));
    enc_scratch_elems = need;
  }
  float* p = enc_scratch;
  auto take = [&](long long n) { float* q = p; p += (n + 63) / 64 * 64; return q; };
  float* audio = take(n_samples);
  float* bufA = take(n_act);      // ELU(x) with history in front (A operand of conv3 / of the strided conv)
  float* bufB = take(n_act);
  float* raw = take(n_act);       // x (skip connection)
  float* hid = take(n_act);       // ELU(conv3(...))  [T][C/2]
  float* xs = take(n_tr);         // transformer residual stream with U replicate rows in front
  float* ln = take(n_tr);
  float* qkvb = take(Tt * 3 * Md);
  float* mid = take(Tt * (long long)c.mimi_ffn);
  float* lat = take((long long)frames * c.ldim);
  float* cond = take((long long)frames * c.d_model);
  PTTS_CUDA_CHECK(cudaMemcpyAsync(audio, audio_any, (size_t)n_samples * sizeof(float), cudaMemcpyDefault, stream));
  int rc;
  // conv0 (1 -> 64, K 7): raw x and ELU(x) behind rk zero history rows
  PTTS_CUDA_CHECK(cudaMemsetAsync(bufA, 0, (size_t)rk * C0 * sizeof(float), stream));
  {
    const long long total = T0 * C0;
    LAUNCH("enc.conv0", (enc_conv0_kernel<<<(unsigned)((total + 255) / 256), 256, 0, stream>>>(audio, n_samples, T0, enc.w0, enc.b0, raw,
                                                                                           bufA + (size_t)rk * C0, C0, c.kernel), cudaGetLastError()));
  }
  long long T = T0;
  int C = C0;
  float* cur = bufA;   // [rk + T][C]: ELU(x), zero history
  float* nxt = bufB;
  for (int i = 0; i < 3; ++i) {
    const EncStageW& S = enc.st[i];
    const int r = S.ratio, hc = C / 2;
    {  // ResBlock conv K3 C -> C/2 on ELU(x), ELU on the way out (seanet.py:27-41)
      GemmEpilogue ep; ep.bias = S.c3_b; ep.act_post = ACT_ELU; ep.out = hid; ep.out_map = plain(hc);
      if ((rc = gemm("enc.res.conv3", cur, false, plain(C), S.c3, (int)T, ep, false)) != PTTS_OK) return rc;
    }
    {  // conv K1 + skip, then the ELU in front of the strided conv; output behind r zero history columns (K - S = r)
      PTTS_CUDA_CHECK(cudaMemsetAsync(nxt, 0, (size_t)r * C * sizeof(float), stream));
      GemmEpilogue ep; ep.bias = S.c1_b; ep.act_post = ACT_ELU; ep.res = raw; ep.res_map = plain(C);
      ep.out = nxt + (size_t)r * C; ep.out_map = plain(C);
      if ((rc = gemm("enc.res.conv1", hid, false, plain(hc), S.c1, (int)T, ep, false)) != PTTS_OK) return rc;
    }
    std::swap(cur, nxt);
    {  // strided conv K = 2r, S = r, C -> 2C: row j = 2r*C elements from j*r*C; raw copy (skip of the next ResBlock) +
       // ELU copy behind the next conv's zero history (rk rows for a ResBlock, last_kernel - 1 for the final conv)
      const long long T2 = T / r;
      const int C2 = 2 * C, nh = (i < 2) ? rk : c.last_kernel - 1;
      PTTS_CUDA_CHECK(cudaMemsetAsync(nxt, 0, (size_t)nh * C2 * sizeof(float), stream));
      GemmEpilogue ep; ep.bias = S.down_b; ep.act_post = ACT_ELU;
      ep.out = nxt + (size_t)nh * C2; ep.out_map = plain(C2);
      if (i < 2) { ep.out2 = raw; ep.out2_map = plain(C2); }
      if ((rc = gemm("enc.down", cur, false, plain(r * C), S.down, (int)T2, ep, false)) != PTTS_OK) return rc;
      std::swap(cur, nxt);
      T = T2; C = C2;
    }
  }
  if (T != Tt) PTTS_FAIL(PTTS_ESTATE, "encoder length bookkeeping (%lld != %lld)", T, Tt);
  float* x = xs + (size_t)U * Md;   // [Tt][Md]
  {  // final conv K3 C -> 512 (seanet.py:103-106)
    GemmEpilogue ep; ep.bias = enc.last_b; ep.out = x; ep.out_map = plain(Md);
    if ((rc = gemm("enc.last", cur, false, plain(C), enc.last, (int)Tt, ep, false)) != PTTS_OK) return rc;
  }
  // ---- encoder transformer (mimi_transformer.py:39-54), whole sequence at once, fp32
  const int NK = c.mimi_context - 1 + ENC_QB;
  const size_t attn_smem = ((((size_t)NK * 65 + 3) & ~(size_t)3) + (size_t)NK * 64 + ENC_QB * 64 + 8 * (size_t)NK) * sizeof(float);
  static DeviceOnce attr_done;
  if (attr_done.first()) PTTS_CUDA_CHECK(cudaFuncSetAttribute(enc_attn_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)attn_smem));
  for (size_t l = 0; l < enc.layers.size(); ++l) {
    const LayerW& L = enc.layers[l];
    LAUNCH("enc.ln", launch_ln_rows(x, L.n1w, L.n1b, 1e-5f, ln, 0, (int)Tt, Md, stream));
    { GemmEpilogue ep; ep.out = qkvb; ep.out_map = plain(3 * Md);
      if ((rc = gemm("enc.qkv", ln, false, plain(Md), L.in_proj, (int)Tt, ep, false)) != PTTS_OK) return rc; }
    {
      dim3 grid((unsigned)((Tt + ENC_QB - 1) / ENC_QB), c.mimi_heads);
      LAUNCH("enc.attn", (enc_attn_kernel<<<grid, 256, attn_smem, stream>>>(qkvb, rope_freqs, ln, (int)Tt, c.mimi_heads, c.mimi_context), cudaGetLastError()));
    }
    { GemmEpilogue ep; ep.out = x; ep.out_map = plain(Md); ep.res = x; ep.res_map = plain(Md); ep.colscale = L.ls1;
      if ((rc = gemm("enc.out_proj", ln, false, plain(Md), L.out_proj, (int)Tt, ep, false)) != PTTS_OK) return rc; }
    LAUNCH("enc.ln", launch_ln_rows(x, L.n2w, L.n2b, 1e-5f, ln, 0, (int)Tt, Md, stream));
    { GemmEpilogue ep; ep.out = mid; ep.out_map = plain(c.mimi_ffn); ep.act_pre = ACT_GELU;
      if ((rc = gemm("enc.ffn1", ln, false, plain(Md), L.lin1, (int)Tt, ep, false)) != PTTS_OK) return rc; }
    { GemmEpilogue ep; ep.out = x; ep.out_map = plain(Md); ep.res = x; ep.res_map = plain(Md); ep.colscale = L.ls2;
      if ((rc = gemm("enc.ffn2", mid, false, plain(c.mimi_ffn), L.lin2, (int)Tt, ep, false)) != PTTS_OK) return rc; }
  }
  // ---- ConvDownsample1d: K = 2U, S = U, no bias, the U history columns replicate the first one (resample.py:20-28)
  LAUNCH("enc.replicate", (enc_replicate_rows_kernel<<<(U * Md + 255) / 256, 256, 0, stream>>>(xs, x, U, Md), cudaGetLastError()));
  { GemmEpilogue ep; ep.out = lat; ep.out_map = plain(c.ldim);
    if ((rc = gemm("enc.downsample", xs, false, plain(U * Md), enc.down, frames, ep, false)) != PTTS_OK) return rc; }
  { GemmEpilogue ep; ep.out = cond; ep.out_map = plain(c.d_model);   // tts_model.py:386-387
    if ((rc = gemm("enc.speaker_proj", lat, false, plain(c.ldim), enc.speaker, frames, ep, false)) != PTTS_OK) return rc; }
  PTTS_CUDA_CHECK(cudaMemcpyAsync(out_cond_any, cond, (size_t)frames * c.d_model * sizeof(float), cudaMemcpyDefault, stream));
  PTTS_CUDA_CHECK(cudaStreamSynchronize(stream));
  return PTTS_OK;
}

}  // namespace ptts
