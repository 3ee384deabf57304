// Mimi-decoder side non-GEMM kernels: latent de-normalisation + 1x1 projection + depthwise
// x16 upsampling (tts_model.py:449-451, resample.py:41-48), streaming-conv history moves
// (conv.py:108-112,155-162 restated as "previous input columns"), the final C->1 conv
// (seanet.py:168-171) and small conversion helpers.
#include "kernels.cuh"

namespace ptts {

// grid (n, C/128): one thread per channel; every weight row a thread needs (32 + 32 floats) is
// fetched with 16-byte loads up front so the loads are in flight together.  `frames` consecutive
// latents of a stream (latents [frames][n][ldim]) are expanded in one launch: frame f uses the
// projection of frame f-1 as its "previous input column" (resample.py:41-48 streaming state).
__global__ void __launch_bounds__(128)
mimi_embed_kernel(const float* __restrict__ latents, const int* __restrict__ slot_ids,
                  const float* __restrict__ emb_std, const float* __restrict__ emb_mean,
                  const float* __restrict__ w_q, const float* __restrict__ w_up,
                  float* __restrict__ e_prev, float* __restrict__ xm, int n, int C, int ldim, int S, int frames) {
  pdl_launch_dependents();
  pdl_wait();
  __shared__ float z[8 * 64];
  const int i = blockIdx.x, slot = slot_ids[i];
  for (int e = threadIdx.x; e < frames * ldim; e += blockDim.x) {
    const int f = e / ldim, j = e - f * ldim;
    z[f * 64 + j] = latents[((long long)f * n + i) * ldim + j] * emb_std[j] + emb_mean[j];
  }
  __syncthreads();
  const int c = blockIdx.y * blockDim.x + threadIdx.x;
  if (c >= C) return;
  const float* w = w_q + (long long)c * ldim;
  const float* wu = w_up + (long long)c * 2 * S;
  float ep = e_prev[(long long)slot * C + c];
  for (int f = 0; f < frames; ++f) {
    float e = 0.0f;
    for (int j = 0; j < ldim; j += 4) {
      const float4 wv = load4(w + j);
      e = fmaf(wv.x, z[f * 64 + j], e);
      e = fmaf(wv.y, z[f * 64 + j + 1], e);
      e = fmaf(wv.z, z[f * 64 + j + 2], e);
      e = fmaf(wv.w, z[f * 64 + j + 3], e);
    }
    // out[t] = e * w[t] + e_prev * w[S + t]   (K = 2S transposed conv, stride S, depthwise)
    for (int t = 0; t < S; t += 4) {
      const float4 a = load4(wu + t), b = load4(wu + S + t);
      float* o = xm + (((long long)i * frames + f) * S + t) * C + c;
      o[0] = e * a.x + ep * b.x;
      o[C] = e * a.y + ep * b.y;
      o[2 * C] = e * a.z + ep * b.z;
      o[3 * C] = e * a.w + ep * b.w;
    }
    ep = e;
  }
  e_prev[(long long)slot * C + c] = ep;
}

cudaError_t launch_mimi_embed(const float* latents, const int* slot_ids, const float* emb_std,
                              const float* emb_mean, const float* w_q, const float* w_up,
                              float* e_prev, float* xm, int n, int C, int ldim, int S, int frames, cudaStream_t st) {
  if (ldim > 64 || ldim % 4 || S % 4 || frames < 1 || frames > 8) return cudaErrorInvalidValue;
  launch_k(mimi_embed_kernel, dim3(n, (C + 127) / 128), 128, 0, st, latents, slot_ids, emb_std, emb_mean, w_q, w_up,
           e_prev, xm, n, C, ldim, S, frames);
  return cudaGetLastError();
}

template <typename TO>
__global__ void stage_rows_kernel(const float* __restrict__ xm, TO* __restrict__ buf, int T, int C,
                                  int hist, long long total4) {
  pdl_launch_dependents();
  pdl_wait();
  const long long q = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (q >= total4) return;
  const long long e = q * 4;
  const int c = (int)(e % C);
  const long long r = e / C;  // i*T + t
  const int t = (int)(r % T);
  const long long i = r / T;
  store4(buf + (i * (hist + T) + hist + t) * C + c, load4(xm + e));
}

cudaError_t launch_stage_rows(const float* xm, void* buf, int buf_type, int n, int T, int C, int hist,
                              cudaStream_t st) {
  const long long total4 = (long long)n * T * C / 4;
  const int blocks = (int)((total4 + 255) / 256);
  if (buf_type == 0) launch_k(stage_rows_kernel<float>, blocks, 256, 0, st, xm, (float*)buf, T, C, hist, total4);
  else launch_k(stage_rows_kernel<bf16>, blocks, 256, 0, st, xm, (bf16*)buf, T, C, hist, total4);
  return cudaGetLastError();
}

// grid (n, count); copies hist*C elements per (stream, conv)
template <typename T, bool GATHER>
__global__ void __launch_bounds__(128)
hist_move_kernel(HistTable tab, const int* __restrict__ slot_ids) {
  pdl_launch_dependents();
  pdl_wait();
  const HistDesc d = tab.d[blockIdx.y];
  const int i = blockIdx.x, slot = slot_ids[i];
  const int total = d.hist * d.C;
  T* state = reinterpret_cast<T*>(d.state) + (long long)slot * total;
  T* buf = reinterpret_cast<T*>(d.buf) + (long long)i * (d.hist + d.T) * d.C;
  if (GATHER) {
    for (int e = threadIdx.x * 4; e < total; e += blockDim.x * 4) store4(buf + e, load4(state + e));
  } else {
    const T* tail = buf + (long long)d.T * d.C;  // rows [T, T+hist) = last `hist` rows of data
    for (int e = threadIdx.x * 4; e < total; e += blockDim.x * 4) store4(state + e, load4(tail + e));
  }
}

cudaError_t launch_hist_gather(const HistTable& t, const int* slot_ids, int type, int n, cudaStream_t st) {
  dim3 grid(n, t.count);
  if (type == 0) launch_k(hist_move_kernel<float, true>, grid, 128, 0, st, t, slot_ids);
  else launch_k(hist_move_kernel<bf16, true>, grid, 128, 0, st, t, slot_ids);
  return cudaGetLastError();
}
cudaError_t launch_hist_scatter(const HistTable& t, const int* slot_ids, int type, int n, cudaStream_t st) {
  dim3 grid(n, t.count);
  if (type == 0) launch_k(hist_move_kernel<float, false>, grid, 128, 0, st, t, slot_ids);
  else launch_k(hist_move_kernel<bf16, false>, grid, 128, 0, st, t, slot_ids);
  return cudaGetLastError();
}

// final conv C -> 1: 128 outputs per CTA, the (128 + K - 1) x C input slab staged in smem
// with a padded row (C + 1) so that thread t reading row t + k is bank-conflict free.
template <typename T>
__global__ void __launch_bounds__(128)
last_conv_kernel(const T* __restrict__ buf, const float* __restrict__ w, const float* __restrict__ b,
                 float* __restrict__ pcm, int Tn, int C, int K) {
  pdl_launch_dependents();
  pdl_wait();
  extern __shared__ float sm[];
  float* sw = sm;               // K*C
  float* sx = sm + K * C;       // (128+K-1) * (C+1)
  const int i = blockIdx.y, t0 = blockIdx.x * 128;
  const int rows = min(128, Tn - t0) + K - 1;
  for (int e = threadIdx.x; e < K * C; e += 128) sw[e] = w[e];
  const T* src = buf + ((long long)i * (Tn + K - 1) + t0) * C;
  for (int e = threadIdx.x * 4; e < rows * C; e += 128 * 4) {
    const float4 v = load4(src + e);
    const int r = e / C, c = e - r * C;
    float* d = sx + r * (C + 1) + c;
    d[0] = v.x; d[1] = v.y; d[2] = v.z; d[3] = v.w;
  }
  __syncthreads();
  const int t = t0 + threadIdx.x;
  if (t >= Tn) return;
  float acc = b[0];
  for (int k = 0; k < K; ++k) {
    const float* xr = sx + (threadIdx.x + k) * (C + 1);
    const float* wr = sw + k * C;
    for (int c = 0; c < C; ++c) acc = fmaf(wr[c], xr[c], acc);
  }
  pcm[(long long)i * Tn + t] = acc;
}

cudaError_t launch_last_conv(const void* buf, int buf_type, const float* w, const float* b, float* pcm,
                             int n, int T, int C, int K, cudaStream_t st) {
  if (C % 4) return cudaErrorInvalidValue;
  dim3 grid((T + 127) / 128, n);
  const size_t smem = (size_t)(K * C + (128 + K - 1) * (C + 1)) * sizeof(float);
  if (smem > 48 * 1024) return cudaErrorInvalidValue;
  if (buf_type == 0) launch_k(last_conv_kernel<float>, grid, 128, smem, st, (const float*)buf, w, b, pcm, T, C, K);
  else launch_k(last_conv_kernel<bf16>, grid, 128, smem, st, (const bf16*)buf, w, b, pcm, T, C, K);
  return cudaGetLastError();
}

__global__ void f32_to_bf16_kernel(const float* __restrict__ s, bf16* __restrict__ d, long long n) {
  pdl_launch_dependents();
  pdl_wait();
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x)
    d[i] = __float2bfloat16_rn(s[i]);
}
cudaError_t launch_f32_to_bf16(const float* src, bf16* dst, long long n, cudaStream_t st) {
  if (n <= 0) return cudaSuccess;
  launch_k(f32_to_bf16_kernel, (int)min(4096LL, (n + 255) / 256), 256, 0, st, src, dst, n);
  return cudaGetLastError();
}
__global__ void fill_f32_kernel(float* __restrict__ d, float v, long long n) {
  pdl_launch_dependents();
  pdl_wait();
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x)
    d[i] = v;
}
cudaError_t launch_fill_f32(float* dst, float v, long long n, cudaStream_t st) {
  if (n <= 0) return cudaSuccess;
  launch_k(fill_f32_kernel, (int)min(4096LL, (n + 255) / 256), 256, 0, st, dst, v, n);
  return cudaGetLastError();
}

}  // namespace ptts

namespace ptts {
__global__ void bf16_to_f32_kernel(const bf16* __restrict__ s, float* __restrict__ d, long long n) {
  pdl_launch_dependents();
  pdl_wait();
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x)
    d[i] = __bfloat162float(s[i]);
}
cudaError_t launch_bf16_to_f32(const bf16* src, float* dst, long long n, cudaStream_t st) {
  if (n <= 0) return cudaSuccess;
  launch_k(bf16_to_f32_kernel, (int)min(4096LL, (n + 255) / 256), 256, 0, st, src, dst, n);
  return cudaGetLastError();
}
__global__ void pcm_to_int16_kernel(const float* __restrict__ pcm, short* __restrict__ out, long long n) {
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) out[i] = (short)(fminf(fmaxf(pcm[i], -1.0f), 1.0f) * 32767.0f);   // C cast truncates, as Tensor.short()
}

cudaError_t launch_pcm_to_int16(const float* pcm, short* out, long long n, cudaStream_t st) {
  if (n <= 0) return cudaSuccess;
  pcm_to_int16_kernel<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(pcm, out, n);
  return cudaGetLastError();
}

}  // namespace ptts
