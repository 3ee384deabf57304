// Engine implementation, part 2: finalize() — resolve the reference's state_dict names
//  into device weights, re-laying convolutions out as GEMM operands.
#include "engine.h"

#include <algorithm>
#include <cstring>

namespace ptts {

// Conv1d weight [C_out, C_in, K] -> GEMM weight [C_out, K*C_in] (k-major, channels-last rows)
static std::vector<float> conv_to_gemm(const HostTensor& w) {
  const int co = (int)w.shape[0], ci = (int)w.shape[1], k = (int)w.shape[2];
  std::vector<float> g((size_t)co * k * ci);
  for (int o = 0; o < co; ++o)
    for (int c = 0; c < ci; ++c)
      for (int t = 0; t < k; ++t) g[((size_t)o * k + t) * ci + c] = w.data[((size_t)o * ci + c) * k + t];
  return g;
}

// ConvTranspose1d weight [C_in, C_out, K=2S] -> [S*C_out, 2*C_in]: row r*C_out+co, column
// j*C_in+ci with j=0 -> tap r+S (multiplies the previous input column), j=1 -> tap r.
// (conv.py:151-163 with K = 2S)
static std::vector<float> convtr_to_gemm(const HostTensor& w, int S) {
  const int ci = (int)w.shape[0], co = (int)w.shape[1], k = (int)w.shape[2];
  std::vector<float> g((size_t)S * co * 2 * ci);
  for (int r = 0; r < S; ++r)
    for (int o = 0; o < co; ++o)
      for (int c = 0; c < ci; ++c) {
        const size_t row = (size_t)r * co + o;
        g[row * 2 * ci + c] = w.data[((size_t)c * co + o) * k + r + S];
        g[row * 2 * ci + ci + c] = w.data[((size_t)c * co + o) * k + r];
      }
  return g;
}

static inline float bf16_round_host(float x) {   // round-to-nearest-even, as __float2bfloat16_rn (finite inputs)
  uint32_t u;
  memcpy(&u, &x, 4);
  u += 0x7FFFu + ((u >> 16) & 1u);
  u &= 0xFFFF0000u;
  float r;
  memcpy(&r, &u, 4);
  return r;
}

// W [N, K], LayerNorm gamma / beta [K]  ->  m = W diag(gamma) (bf16 copy is what the GEMM reads),
// c1[n] = sum_k bf16(W[n,k] gamma[k]),  c2[n] = sum_k W[n,k] beta[k]      
int Engine::make_ln_folded(const std::string& wname, const std::string& gname, const std::string& bname, int N, int K,
                           Mat* m, float** c1, float** c2) {
  auto iw = host_w.find(wname), ig = host_w.find(gname), ib = host_w.find(bname);
  if (iw == host_w.end() || ig == host_w.end() || ib == host_w.end() || iw->second.numel() != (int64_t)N * K ||
      ig->second.numel() != K || ib->second.numel() != K)
    PTTS_FAIL(PTTS_EINVAL, "missing/malformed tensors for the folded LayerNorm of %s", wname.c_str());
  const std::vector<float>&W = iw->second.data, &G = ig->second.data, &Bt = ib->second.data;
  std::vector<float> wf((size_t)N * K), v1(N), v2(N);
  for (int n = 0; n < N; ++n) {
    double s1 = 0.0, s2 = 0.0;
    for (int k = 0; k < K; ++k) {
      const float w = W[(size_t)n * K + k], wg = w * G[k];
      wf[(size_t)n * K + k] = wg;
      s1 += (double)bf16_round_host(wg);
      s2 += (double)w * (double)Bt[k];
    }
    v1[n] = (float)s1; v2[n] = (float)s2;
  }
  int rc;
  if ((rc = make_mat_from(wf, N, K, m)) != PTTS_OK) return rc;
  *c1 = upload(v1);
  *c2 = upload(v2);
  if (!*c1 || !*c2) PTTS_FAIL(PTTS_ECUDA, "upload failed");
  return PTTS_OK;
}

int Engine::finalize() {
  if (ready) return PTTS_OK;
  const ptts_config& c = cfg;
  const int64_t D = c.d_model, F = c.flow_dim, L = c.ldim, Md = c.mimi_d;
  int rc;
#define GETV(dst, name, ...)                                   \
  do {                                                         \
    dst = dev_f32(name, std::vector<int64_t>{__VA_ARGS__});    \
    if (!dst) return PTTS_EINVAL;                              \
  } while (0)
#define GETM(mat, name, N, K, ...)                                         \
  do {                                                                     \
    if ((rc = make_mat(name, (int)(N), (int)(K), &mat, ##__VA_ARGS__)) != PTTS_OK) return rc; \
  } while (0)

  GETV(bos_emb, "flow_lm.bos_emb", L);
  GETV(w_in, "flow_lm.input_linear.weight", D, L);
  GETV(out_norm_w, "flow_lm.out_norm.weight", D);
  GETV(out_norm_b, "flow_lm.out_norm.bias", D);
  GETV(eos_w, "flow_lm.out_eos.weight", 1, D);
  GETV(eos_b, "flow_lm.out_eos.bias", 1);
  GETV(emb_std, "flow_lm.emb_std", L);
  GETV(emb_mean, "flow_lm.emb_mean", L);
  GETV(rope_freqs, "rope.freqs", 32);
  {
    auto it = host_w.find("flow.tcomb");
    if (it == host_w.end() || it->second.shape.size() != 2 || it->second.shape[1] != F)
      PTTS_FAIL(PTTS_EINVAL, "missing or malformed flow.tcomb [lsd_steps, flow_dim]");
    tcomb_rows = (int)it->second.shape[0];
    tcomb = upload(it->second.data);
    if (!tcomb) PTTS_FAIL(PTTS_ECUDA, "upload failed");
  }
  {  // optional: the text conditioner's embedding table, for prefill from token ids
    auto it = host_w.find("flow_lm.conditioner.embed.weight");
    if (it != host_w.end()) {
      if (it->second.shape.size() != 2 || it->second.shape[1] != D)
        PTTS_FAIL(PTTS_EINVAL, "flow_lm.conditioner.embed.weight must be [rows, d_model]");
      text_embed_rows = (int)it->second.shape[0];
      text_embed = upload(it->second.data);
      if (!text_embed) PTTS_FAIL(PTTS_ECUDA, "upload failed");
    }
  }
  flow_layers.resize(c.n_layers);
  for (int i = 0; i < c.n_layers; ++i) {
    const std::string p = "flow_lm.transformer.layers." + std::to_string(i) + ".";
    LayerW& w = flow_layers[i];
    GETV(w.n1w, p + "norm1.weight", D);
    GETV(w.n1b, p + "norm1.bias", D);
    GETV(w.n2w, p + "norm2.weight", D);
    GETV(w.n2b, p + "norm2.bias", D);
    GETM(w.in_proj, p + "self_attn.in_proj.weight", 3 * D, D);
    GETM(w.out_proj, p + "self_attn.out_proj.weight", D, D);
    GETM(w.lin1, p + "linear1.weight", c.ffn, D);
    GETM(w.lin2, p + "linear2.weight", D, c.ffn);
    if (act_type == 1) {
      if ((rc = make_ln_folded(p + "self_attn.in_proj.weight", p + "norm1.weight", p + "norm1.bias", 3 * (int)D, (int)D,
                               &w.in_proj_ln, &w.qkv_c1, &w.qkv_c2)) != PTTS_OK) return rc;
      if ((rc = make_ln_folded(p + "linear1.weight", p + "norm2.weight", p + "norm2.bias", c.ffn, (int)D, &w.lin1_ln,
                               &w.ffn_c1, &w.ffn_c2)) != PTTS_OK) return rc;
    }
  }
  // ---- flow head (mlp.py:134-215)
  const std::string fp = "flow_lm.flow_net.";
  GETM(cond_embed, fp + "cond_embed.weight", F, D);
  GETV(cond_b, fp + "cond_embed.bias", F);
  {
    std::vector<float> bt(host_w[fp + "cond_embed.bias"].data);
    const std::vector<float>& tc = host_w["flow.tcomb"].data;
    for (int64_t j = 0; j < F; ++j) bt[j] += tc[j];
    cond_b_t0 = upload(bt);
    if (!cond_b_t0) PTTS_FAIL(PTTS_ECUDA, "upload failed");
  }
  GETM(input_proj, fp + "input_proj.weight", F, L, false);  // A operand is the fp32 latent
  GETV(inproj_b, fp + "input_proj.bias", F);
  GETM(final_lin, fp + "final_layer.linear.weight", L, F);
  GETV(final_b, fp + "final_layer.linear.bias", L);
  flow_blocks.resize(c.flow_depth);
  std::vector<float> modw, modb;  // all adaLN modulations concatenated: [depth*3F + 2F, F]
  for (int j = 0; j < c.flow_depth; ++j) {
    const std::string p = fp + "res_blocks." + std::to_string(j) + ".";
    FlowBlockW& w = flow_blocks[j];
    GETV(w.lnw, p + "in_ln.weight", F);
    GETV(w.lnb, p + "in_ln.bias", F);
    GETM(w.mlp0, p + "mlp.0.weight", F, F);
    GETV(w.b0, p + "mlp.0.bias", F);
    GETM(w.mlp2, p + "mlp.2.weight", F, F);
    GETV(w.b2, p + "mlp.2.bias", F);
    auto iw = host_w.find(p + "adaLN_modulation.1.weight"), ib = host_w.find(p + "adaLN_modulation.1.bias");
    if (iw == host_w.end() || ib == host_w.end() || iw->second.numel() != 3 * F * F || ib->second.numel() != 3 * F)
      PTTS_FAIL(PTTS_EINVAL, "missing/malformed %sadaLN_modulation.1", p.c_str());
    modw.insert(modw.end(), iw->second.data.begin(), iw->second.data.end());
    modb.insert(modb.end(), ib->second.data.begin(), ib->second.data.end());
  }
  {
    auto iw = host_w.find(fp + "final_layer.adaLN_modulation.1.weight");
    auto ib = host_w.find(fp + "final_layer.adaLN_modulation.1.bias");
    if (iw == host_w.end() || ib == host_w.end() || iw->second.numel() != 2 * F * F || ib->second.numel() != 2 * F)
      PTTS_FAIL(PTTS_EINVAL, "missing/malformed final_layer.adaLN_modulation.1");
    modw.insert(modw.end(), iw->second.data.begin(), iw->second.data.end());
    modb.insert(modb.end(), ib->second.data.begin(), ib->second.data.end());
  }
  if ((rc = make_mat_from(modw, (int)((c.flow_depth * 3 + 2) * F), (int)F, &mod_all)) != PTTS_OK) return rc;
  mod_b = upload(modb);

  // ---- mimi
  GETV(wq, "mimi.quantizer.output_proj.weight", Md, L, 1);
  GETV(wup, "mimi.upsample.convtr.convtr.weight", Md, 1, 2 * c.upsample_stride);
  mimi_layers.resize(c.mimi_layers);
  for (int i = 0; i < c.mimi_layers; ++i) {
    const std::string p = "mimi.decoder_transformer.transformer.layers." + std::to_string(i) + ".";
    LayerW& w = mimi_layers[i];
    GETV(w.n1w, p + "norm1.weight", Md);
    GETV(w.n1b, p + "norm1.bias", Md);
    GETV(w.n2w, p + "norm2.weight", Md);
    GETV(w.n2b, p + "norm2.bias", Md);
    GETV(w.ls1, p + "layer_scale_1.scale", Md);
    GETV(w.ls2, p + "layer_scale_2.scale", Md);
    GETM(w.in_proj, p + "self_attn.in_proj.weight", 3 * Md, Md);
    GETM(w.out_proj, p + "self_attn.out_proj.weight", Md, Md);
    GETM(w.lin1, p + "linear1.weight", c.mimi_ffn, Md);
    GETM(w.lin2, p + "linear2.weight", Md, c.mimi_ffn);
  }
  auto need = [&](const std::string& n) -> const HostTensor* {
    auto it = host_w.find(n);
    if (it == host_w.end()) { set_error("missing tensor " + n); return nullptr; }
    return &it->second;
  };
  int C = c.n_filters * 8;
  {
    const HostTensor* w = need("mimi.decoder.model.0.conv.weight");
    if (!w) return PTTS_EINVAL;
    if (w->shape != std::vector<int64_t>{C, Md, c.kernel}) PTTS_FAIL(PTTS_EINVAL, "bad decoder.model.0 shape");
    if ((rc = make_mat_from(conv_to_gemm(*w), C, c.kernel * (int)Md, &conv0)) != PTTS_OK) return rc;
    GETV(conv0_b, "mimi.decoder.model.0.conv.bias", C);
  }
  int idx = 1;
  for (int i = 0; i < 3; ++i) {
    const int r = c.ratios[i], C2 = C / 2, hc = C2 / 2;
    const std::string pt = "mimi.decoder.model." + std::to_string(idx + 1) + ".convtr.";
    const std::string pr = "mimi.decoder.model." + std::to_string(idx + 2) + ".block.";
    const HostTensor* w = need(pt + "weight");
    const HostTensor* b = need(pt + "bias");
    if (!w || !b) return PTTS_EINVAL;
    if (w->shape != std::vector<int64_t>{C, C2, 2 * r} || b->numel() != C2)
      PTTS_FAIL(PTTS_EINVAL, "bad shape for %sweight", pt.c_str());
    if ((rc = make_mat_from(convtr_to_gemm(*w, r), r * C2, 2 * C, &stages[i].tr)) != PTTS_OK) return rc;
    std::vector<float> be((size_t)r * C2);
    for (int q = 0; q < r; ++q)
      for (int o = 0; o < C2; ++o) be[(size_t)q * C2 + o] = b->data[o];
    stages[i].tr_b = upload(be);
    const HostTensor* w3 = need(pr + "1.conv.weight");
    const HostTensor* w1 = need(pr + "3.conv.weight");
    if (!w3 || !w1) return PTTS_EINVAL;
    if (w3->shape != std::vector<int64_t>{hc, C2, c.res_kernel} || w1->shape != std::vector<int64_t>{C2, hc, 1})
      PTTS_FAIL(PTTS_EINVAL, "bad resblock shapes at %s", pr.c_str());
    // bf16 mode: the hidden width is zero-padded to a multiple of 64 channels so that the 1x1 conv's
    // reduction dim fills a 128-byte swizzle row (tcgen05 path); padded channels stay exactly 0:
    // zero weight rows + zero bias -> ELU(0) = 0 -> multiplied by zero weight columns.
    const int hp = hid_width(hc);
    {
      std::vector<float> g3 = conv_to_gemm(*w3), g1 = conv_to_gemm(*w1);
      const int k3 = c.res_kernel * C2;
      std::vector<float> p3((size_t)hp * k3, 0.0f), p1((size_t)C2 * hp, 0.0f), b3(hp, 0.0f);
      for (int o = 0; o < hc; ++o) std::copy(g3.begin() + (size_t)o * k3, g3.begin() + (size_t)(o + 1) * k3, p3.begin() + (size_t)o * k3);
      for (int o = 0; o < C2; ++o) std::copy(g1.begin() + (size_t)o * hc, g1.begin() + (size_t)(o + 1) * hc, p1.begin() + (size_t)o * hp);
      const HostTensor* bb = need(pr + "1.conv.bias");
      if (!bb || bb->numel() != hc) PTTS_FAIL(PTTS_EINVAL, "bad bias for %s1.conv", pr.c_str());
      std::copy(bb->data.begin(), bb->data.end(), b3.begin());
      if ((rc = make_mat_from(p3, hp, k3, &stages[i].c3)) != PTTS_OK) return rc;
      if ((rc = make_mat_from(p1, C2, hp, &stages[i].c1)) != PTTS_OK) return rc;
      stages[i].c3_b = upload(b3);
      res_b3_host[i] = b3;
    }
    GETV(stages[i].c1_b, pr + "3.conv.bias", C2);
    res_b1_host[i] = host_w[pr + "3.conv.bias"].data;
    C = C2;
    idx += 3;
  }
  {
    const std::string pl = "mimi.decoder.model." + std::to_string(idx + 1) + ".conv.";
    const HostTensor* w = need(pl + "weight");
    if (!w) return PTTS_EINVAL;
    if (w->shape != std::vector<int64_t>{1, C, c.last_kernel}) PTTS_FAIL(PTTS_EINVAL, "bad last conv shape");
    last_w = upload(conv_to_gemm(*w));
    GETV(last_b, pl + "bias", 1);
    // host copies of the last stage's per-channel vectors for the fused stage kernel (seanet_fused.cu)
    if (C == 64 && c.last_kernel == 3 && c.ratios[2] == 4) {
      const std::vector<float> lw = conv_to_gemm(*w);
      const std::string pt2 = "mimi.decoder.model." + std::to_string(idx - 2) + ".convtr.bias";
      const std::string pr2 = "mimi.decoder.model." + std::to_string(idx - 1) + ".block.";
      const HostTensor *bt = need(pt2), *b3 = need(pr2 + "1.conv.bias"), *b1 = need(pr2 + "3.conv.bias"), *bl = need(pl + "bias");
      if (!bt || !b3 || !b1 || !bl) return PTTS_EINVAL;
      if (bt->numel() == 64 && b3->numel() == 32 && b1->numel() == 64 && lw.size() == 192) {
        for (int j = 0; j < 256; ++j) fused2_consts.b_tr[j] = bt->data[j & 63];
        for (int j = 0; j < 64; ++j) fused2_consts.b_c1[j] = b1->data[j];
        for (int j = 0; j < 32; ++j) fused2_consts.b_c3[j] = b3->data[j];
        for (int j = 0; j < 192; ++j) fused2_consts.w_last[j] = lw[j];
        fused2_consts.b_last = bl->data[0];
      } else {
        use_fused_stage2 = false;
      }
    } else {
      use_fused_stage2 = false;
    }
  }
  if ((rc = load_encoder()) != PTTS_OK) return rc;   // optional: mimi.encoder* / mimi.downsample* / speaker_proj_weight
  PTTS_CUDA_CHECK(cudaStreamSynchronize(stream));
  host_w.clear();
  ready = true;
  return PTTS_OK;
#undef GETV
#undef GETM
}

}  // namespace ptts
