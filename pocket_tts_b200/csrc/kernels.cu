// Non-GEMM kernels of the Pocket TTS frame step (sm_100a).  All are bandwidth / latency
// kernels: coalesced, vectorised, warp-shuffle reductions (north_star: "KV-cache attention,
// norms, RoPE, ELU, the flow-head epilogue and state-buffer shifts").
#include "kernels.cuh"

namespace ptts {

PTTS_TL_ENABLE_FN(kernels)

// =====================================================================================
// FlowLM input embedding                                              flow_lm.py:121-122
// =====================================================================================
__global__ void __launch_bounds__(256)
flow_embed_kernel(const float* __restrict__ prev, const int* __restrict__ slot_ids,
                  const float* __restrict__ bos, const float* __restrict__ w_in,
                  float* __restrict__ h, int d, int ldim) {
  pdl_launch_dependents();
  const unsigned tl = tl_begin(5);
  // input_linear's weight (128 KB, cold every step) -> L2 while the previous kernel is still running
  for (int off = (blockIdx.x * blockDim.x + threadIdx.x) * 128; off < d * ldim * 4; off += gridDim.x * blockDim.x * 128)
    asm volatile("prefetch.global.L2 [%0];" ::"l"(reinterpret_cast<const char*>(w_in) + off));
  pdl_wait();
  tl_mark(tl, 2);
  __shared__ float seq[64];
  const int i = blockIdx.x, slot = slot_ids[i];
  if (threadIdx.x < ldim) {
    const float v = prev[slot * ldim + threadIdx.x];
    seq[threadIdx.x] = isnan(v) ? bos[threadIdx.x] : v;  // NaN marks BOS (tts_model.py:748-753)
  }
  __syncthreads();
  for (int c = threadIdx.x; c < d; c += blockDim.x) {
    const float* w = w_in + (long long)c * ldim;
    float acc = 0.0f;
    for (int j = 0; j < ldim; j += 4) {
      const float4 wv = load4(w + j);
      acc = fmaf(wv.x, seq[j], acc);
      acc = fmaf(wv.y, seq[j + 1], acc);
      acc = fmaf(wv.z, seq[j + 2], acc);
      acc = fmaf(wv.w, seq[j + 3], acc);
    }
    h[(long long)i * d + c] = acc;
  }
}

cudaError_t launch_flow_embed(const float* prev_latent, const int* slot_ids, const float* bos,
                              const float* w_in, float* h, int n, int d, int ldim, cudaStream_t st) {
  if (ldim > 64 || ldim % 4) return cudaErrorInvalidValue;
  launch_k(flow_embed_kernel, n, 256, 0, st, prev_latent, slot_ids, bos, w_in, h, d, ldim);
  return cudaGetLastError();
}

__global__ void embed_gather_kernel(const int* __restrict__ tokens, const float* __restrict__ table,
                                    float* __restrict__ h, int d) {
  pdl_launch_dependents();
  pdl_wait();
  const float* src = table + (long long)tokens[blockIdx.x] * d;
  float* dst = h + (long long)blockIdx.x * d;
  for (int c = threadIdx.x * 4; c < d; c += blockDim.x * 4) store4(dst + c, load4(src + c));
}

cudaError_t launch_embed_gather(const int* tokens, const float* table, float* h, int rows, int d, cudaStream_t st) {
  if (d % 4) return cudaErrorInvalidValue;
  launch_k(embed_gather_kernel, rows, 256, 0, st, tokens, table, h, d);
  return cudaGetLastError();
}

// =====================================================================================
// LayerNorm over rows (nn.LayerNorm, biased variance)      mimi_transformer.py:26-27,47,41
// one warp-group of 128 threads per row, d <= 1024: each thread owns 2 float4
// =====================================================================================
template <typename TO, int VPT>
__global__ void __launch_bounds__(128)
ln_rows_kernel(const float* __restrict__ x, const float* __restrict__ w, const float* __restrict__ b,
               float eps, TO* __restrict__ y, int d) {
  pdl_launch_dependents();
  const unsigned tl = tl_begin(3);
  pdl_wait();
  tl_mark(tl, 2);
  __shared__ float scratch[33];
  const long long row = blockIdx.x;
  const float* xr = x + row * d;
  float4 v[VPT];
  float s = 0.0f;
#pragma unroll
  for (int j = 0; j < VPT; ++j) {
    const int c = (threadIdx.x + j * 128) * 4;
    v[j] = (c < d) ? load4(xr + c) : make_float4(0, 0, 0, 0);
    s += v[j].x + v[j].y + v[j].z + v[j].w;
  }
  const float mu = block_sum(s, scratch) / (float)d;
  float q = 0.0f;
#pragma unroll
  for (int j = 0; j < VPT; ++j) {
    const int c = (threadIdx.x + j * 128) * 4;
    if (c < d) {
      const float a0 = v[j].x - mu, a1 = v[j].y - mu, a2 = v[j].z - mu, a3 = v[j].w - mu;
      q += a0 * a0 + a1 * a1 + a2 * a2 + a3 * a3;
    }
  }
  const float rstd = rsqrtf(block_sum(q, scratch) / (float)d + eps);
#pragma unroll
  for (int j = 0; j < VPT; ++j) {
    const int c = (threadIdx.x + j * 128) * 4;
    if (c < d) {
      const float4 wv = load4(w + c), bv = load4(b + c);
      float4 o;
      o.x = (v[j].x - mu) * rstd * wv.x + bv.x;
      o.y = (v[j].y - mu) * rstd * wv.y + bv.y;
      o.z = (v[j].z - mu) * rstd * wv.z + bv.z;
      o.w = (v[j].w - mu) * rstd * wv.w + bv.w;
      store4(y + row * d + c, o);
    }
  }
}

// d <= 512 (Mimi transformers: thousands of rows per pass): one WARP per row, 8 rows per CTA, no block barriers;
// the choice depends on d only, so a row is normalised by the same arithmetic whatever the pass size
template <typename TO>
__global__ void __launch_bounds__(256)
ln_rows_warp_kernel(const float* __restrict__ x, const float* __restrict__ w, const float* __restrict__ b, float eps,
                    TO* __restrict__ y, int rows, int d) {
  pdl_launch_dependents();
  pdl_wait();
  const int lane = threadIdx.x & 31;
  const long long row = (long long)blockIdx.x * 8 + (threadIdx.x >> 5);
  if (row >= rows) return;
  const float* xr = x + row * d;
  float4 v[4];
  float s = 0.0f;
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    const int c = (lane + j * 32) * 4;
    v[j] = (c < d) ? load4(xr + c) : make_float4(0, 0, 0, 0);
    s += (v[j].x + v[j].y) + (v[j].z + v[j].w);
  }
  const float mu = warp_sum(s) / (float)d;
  float q = 0.0f;
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    if ((lane + j * 32) * 4 < d) {
      const float a0 = v[j].x - mu, a1 = v[j].y - mu, a2 = v[j].z - mu, a3 = v[j].w - mu;
      q += (a0 * a0 + a1 * a1) + (a2 * a2 + a3 * a3);
    }
  }
  const float rstd = rsqrtf(warp_sum(q) / (float)d + eps);
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    const int c = (lane + j * 32) * 4;
    if (c < d) {
      const float4 wv = load4(w + c), bv = load4(b + c);
      store4(y + row * d + c, make_float4((v[j].x - mu) * rstd * wv.x + bv.x, (v[j].y - mu) * rstd * wv.y + bv.y,
                                          (v[j].z - mu) * rstd * wv.z + bv.z, (v[j].w - mu) * rstd * wv.w + bv.w));
    }
  }
}

cudaError_t launch_ln_rows(const float* x, const float* w, const float* b, float eps, void* y,
                           int y_type, int rows, int d, cudaStream_t st) {
  if (rows <= 0) return cudaSuccess;
  if (d % 4 || d > 1024) return cudaErrorInvalidValue;
  if (d <= 512) {
    if (y_type == 0) launch_k(ln_rows_warp_kernel<float>, (rows + 7) / 8, 256, 0, st, x, w, b, eps, (float*)y, rows, d);
    else launch_k(ln_rows_warp_kernel<bf16>, (rows + 7) / 8, 256, 0, st, x, w, b, eps, (bf16*)y, rows, d);
    return cudaGetLastError();
  }
  if (y_type == 0) launch_k(ln_rows_kernel<float, 2>, rows, 128, 0, st, x, w, b, eps, (float*)y, d);
  else launch_k(ln_rows_kernel<bf16, 2>, rows, 128, 0, st, x, w, b, eps, (bf16*)y, d);
  return cudaGetLastError();
}

// =====================================================================================
// out_norm + EOS head                                          flow_lm.py:128-129,153-154
// =====================================================================================
template <typename TO>
__global__ void __launch_bounds__(128)
final_norm_eos_kernel(const float* __restrict__ h, const float* __restrict__ w,
                      const float* __restrict__ b, float eps, const float* __restrict__ w_eos,
                      const float* __restrict__ b_eos, TO* __restrict__ c,
                      float* __restrict__ eos_logit, int d, int* __restrict__ offsets,
                      const int* __restrict__ slot_ids) {
  pdl_launch_dependents();
  const unsigned tl = tl_begin(4);
  pdl_wait();
  tl_mark(tl, 2);
  __shared__ float scratch[33];
  const long long row = blockIdx.x;
  const float* xr = h + row * d;
  float4 v[2];
  float s = 0.0f;
#pragma unroll
  for (int j = 0; j < 2; ++j) {
    const int col = (threadIdx.x + j * 128) * 4;
    v[j] = (col < d) ? load4(xr + col) : make_float4(0, 0, 0, 0);
    s += v[j].x + v[j].y + v[j].z + v[j].w;
  }
  const float mu = block_sum(s, scratch) / (float)d;
  float q = 0.0f;
#pragma unroll
  for (int j = 0; j < 2; ++j) {
    const int col = (threadIdx.x + j * 128) * 4;
    if (col < d) {
      const float a0 = v[j].x - mu, a1 = v[j].y - mu, a2 = v[j].z - mu, a3 = v[j].w - mu;
      q += a0 * a0 + a1 * a1 + a2 * a2 + a3 * a3;
    }
  }
  const float rstd = rsqrtf(block_sum(q, scratch) / (float)d + eps);
  float dot = 0.0f;
#pragma unroll
  for (int j = 0; j < 2; ++j) {
    const int col = (threadIdx.x + j * 128) * 4;
    if (col < d) {
      const float4 wv = load4(w + col), bv = load4(b + col), ev = load4(w_eos + col);
      float4 o;
      o.x = (v[j].x - mu) * rstd * wv.x + bv.x;
      o.y = (v[j].y - mu) * rstd * wv.y + bv.y;
      o.z = (v[j].z - mu) * rstd * wv.z + bv.z;
      o.w = (v[j].w - mu) * rstd * wv.w + bv.w;
      dot += o.x * ev.x + o.y * ev.y + o.z * ev.z + o.w * ev.w;
      store4(c + row * d + col, o);
    }
  }
  dot = block_sum(dot, scratch);
  if (threadIdx.x == 0) {
    eos_logit[row] = dot + b_eos[0];
    // the step's attention kernels have all read the stream's offset: advance it (tts_model.py:342-345)
    if (offsets) offsets[slot_ids[row]] += 1;
  }
}

cudaError_t launch_final_norm_eos(const float* h, const float* w, const float* b, float eps,
                                  const float* w_eos, const float* b_eos, void* c, int c_type,
                                  float* eos_logit, int n, int d, cudaStream_t st, int* offsets,
                                  const int* slot_ids) {
  if (d % 4 || d > 1024) return cudaErrorInvalidValue;
  if (c_type == 0)
    launch_k(final_norm_eos_kernel<float>, n, 128, 0, st, h, w, b, eps, w_eos, b_eos, (float*)c, eos_logit, d, offsets, slot_ids);
  else
    launch_k(final_norm_eos_kernel<bf16>, n, 128, 0, st, h, w, b, eps, w_eos, b_eos, (bf16*)c, eos_logit, d, offsets, slot_ids);
  return cudaGetLastError();
}

// =====================================================================================
// flow-head elementwise kernels                                        mlp.py:16-17,49-55
// =====================================================================================
template <typename TO>
__global__ void silu_add_kernel(const float* __restrict__ cemb, const float* __restrict__ trow,
                                TO* __restrict__ sy, long long total, int d) {
  pdl_launch_dependents();
  pdl_wait();
  const long long i = ((long long)blockIdx.x * blockDim.x + threadIdx.x) * 4;
  if (i >= total) return;
  const int c = (int)(i % d);
  const float4 a = load4(cemb + i), t = load4(trow + c);
  store4(sy + i, make_float4(silu_f(a.x + t.x), silu_f(a.y + t.y), silu_f(a.z + t.z), silu_f(a.w + t.w)));
}

cudaError_t launch_silu_add(const float* cemb, const float* tcomb_row, void* sy, int sy_type, int n,
                            int d, cudaStream_t st) {
  const long long total = (long long)n * d;
  const int blocks = (int)((total / 4 + 255) / 256);
  if (sy_type == 0) launch_k(silu_add_kernel<float>, blocks, 256, 0, st, cemb, tcomb_row, (float*)sy, total, d);
  else launch_k(silu_add_kernel<bf16>, blocks, 256, 0, st, cemb, tcomb_row, (bf16*)sy, total, d);
  return cudaGetLastError();
}

// custom LayerNorm of the flow head: (x - mean) / sqrt(var + eps), biased var (mlp.py:49-55),
// followed by modulate(x, shift, scale) = x * (1 + scale) + shift (mlp.py:16-17)
template <typename TO>
__global__ void __launch_bounds__(128)
ln_modulate_kernel(const float* __restrict__ u, const float* __restrict__ w,
                   const float* __restrict__ b, const float* __restrict__ shift,
                   const float* __restrict__ scale, int mod_ld, TO* __restrict__ g, int d) {
  pdl_launch_dependents();
  pdl_wait();
  __shared__ float scratch[33];
  const long long row = blockIdx.x;
  const int c = threadIdx.x * 4;
  const bool on = c < d;
  const float4 v = on ? load4(u + row * d + c) : make_float4(0, 0, 0, 0);
  const float mu = block_sum(v.x + v.y + v.z + v.w, scratch) / (float)d;
  const float a0 = v.x - mu, a1 = v.y - mu, a2 = v.z - mu, a3 = v.w - mu;
  const float var = block_sum(on ? (a0 * a0 + a1 * a1 + a2 * a2 + a3 * a3) : 0.0f, scratch) / (float)d;
  const float den = sqrtf(var + 1e-6f);
  if (!on) return;
  float4 o = make_float4(a0 / den, a1 / den, a2 / den, a3 / den);
  if (w) {
    const float4 wv = load4(w + c), bv = load4(b + c);
    o.x = o.x * wv.x + bv.x; o.y = o.y * wv.y + bv.y; o.z = o.z * wv.z + bv.z; o.w = o.w * wv.w + bv.w;
  }
  const float4 sh = load4(shift + row * mod_ld + c), sc = load4(scale + row * mod_ld + c);
  o.x = o.x * (1.0f + sc.x) + sh.x;
  o.y = o.y * (1.0f + sc.y) + sh.y;
  o.z = o.z * (1.0f + sc.z) + sh.z;
  o.w = o.w * (1.0f + sc.w) + sh.w;
  store4(g + row * d + c, o);
}

cudaError_t launch_ln_modulate(const float* u, const float* w, const float* b, const float* shift,
                               const float* scale, int mod_ld, void* g, int g_type, int n, int d,
                               cudaStream_t st) {
  if (d > 512 || d % 4) return cudaErrorInvalidValue;
  if (g_type == 0)
    launch_k(ln_modulate_kernel<float>, n, 128, 0, st, u, w, b, shift, scale, mod_ld, (float*)g, d);
  else
    launch_k(ln_modulate_kernel<bf16>, n, 128, 0, st, u, w, b, shift, scale, mod_ld, (bf16*)g, d);
  return cudaGetLastError();
}

__global__ void commit_latent_kernel(const float* __restrict__ x, const int* __restrict__ slot_ids,
                                     float* __restrict__ prev, float* __restrict__ out, int n, int ldim) {
  pdl_launch_dependents();
  pdl_wait();
  const int idx = blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= n * ldim) return;
  const int i = idx / ldim, j = idx - i * ldim;
  const float v = x[idx];
  prev[slot_ids[i] * ldim + j] = v;
  if (out) out[idx] = v;
}

cudaError_t launch_commit_latent(const float* x, const int* slot_ids, float* prev_latent, float* out,
                                 int n, int ldim, cudaStream_t st) {
  launch_k(commit_latent_kernel, (n * ldim + 255) / 256, 256, 0, st, x, slot_ids, prev_latent, out, n, ldim);
  return cudaGetLastError();
}

__global__ void advance_kernel(int* __restrict__ offsets, const int* __restrict__ slot_ids,
                               const int* __restrict__ inc, int const_inc, int n) {
  pdl_launch_dependents();
  pdl_wait();
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) offsets[slot_ids[i]] += inc ? inc[i] : const_inc;
}

cudaError_t launch_advance(int* offsets, const int* slot_ids, const int* inc, int const_inc, int n,
                           cudaStream_t st) {
  launch_k(advance_kernel, (n + 127) / 128, 128, 0, st, offsets, slot_ids, inc, const_inc, n);
  return cudaGetLastError();
}

}  // namespace ptts
