// FP32-accurate SIMT GEMM (FFMA, fp32 accumulate).  This is the parity-mode back-end
// (PTTS_PRECISION_FP32: latent rel-L2 <= 1e-4 needs >tf32 precision)
// and the numerical cross-check for the tcgen05 back-end.  Classic 2-stage register-prefetch
// tiling: BMxBNx16 tiles, 256 threads, 4x4 outputs per thread, operands transposed into
// shared memory so the inner product reads are float4 and conflict free.
#include "gemm.cuh"

namespace ptts {

template <typename TA, typename TO, int BM, int BN>
__global__ void __launch_bounds__(256)
gemm_simt_kernel(const TA* __restrict__ A, RowMap amap, const TA* __restrict__ W, int M, int N, int K,
                 GemmEpilogue ep) {
  pdl_launch_dependents();
  pdl_wait();
  constexpr int BK = 16;
  constexpr int TX = BN / 4;          // threads along N
  constexpr int LA = (BM * 4 + 255) / 256;  // float4 loads of A per thread per k-tile
  constexpr int LW = (BN * 4 + 255) / 256;
  static_assert((BM / 4) * (BN / 4) == 256, "tile/thread mismatch");
  __shared__ __align__(16) float As[2][BK][BM + 4];
  __shared__ __align__(16) float Ws[2][BK][BN + 4];

  const int tid = threadIdx.x;
  const int tx = tid % TX, ty = tid / TX;
  const int m0 = blockIdx.y * BM, n0 = blockIdx.x * BN;

  // per-thread global load coordinates
  const TA* a_ptr[LA];
  int a_r[LA], a_kq[LA];
#pragma unroll
  for (int l = 0; l < LA; ++l) {
    const int idx = tid + l * 256;
    const int r = idx >> 2, kq = idx & 3;
    a_r[l] = r; a_kq[l] = kq;
    const int m = m0 + r;
    const TA* p = nullptr;
    if (idx < BM * 4 && m < M) {
      const int b = m / amap.rows_per_batch, t = m - b * amap.rows_per_batch;
      if (t < amap.valid_rows) p = A + b * amap.batch_stride + (long long)t * amap.ld + kq * 4;
    }
    a_ptr[l] = p;
  }
  const TA* w_ptr[LW];
  int w_r[LW], w_kq[LW];
#pragma unroll
  for (int l = 0; l < LW; ++l) {
    const int idx = tid + l * 256;
    const int r = idx >> 2, kq = idx & 3;
    w_r[l] = r; w_kq[l] = kq;
    const int n = n0 + r;
    w_ptr[l] = (idx < BN * 4 && n < N) ? W + (long long)n * K + kq * 4 : nullptr;
  }

  float4 ra[LA], rw[LW];
  auto gload = [&](int k0) {
#pragma unroll
    for (int l = 0; l < LA; ++l) ra[l] = a_ptr[l] ? load4(a_ptr[l] + k0) : make_float4(0, 0, 0, 0);
#pragma unroll
    for (int l = 0; l < LW; ++l) rw[l] = w_ptr[l] ? load4(w_ptr[l] + k0) : make_float4(0, 0, 0, 0);
  };
  auto sstore = [&](int buf) {
#pragma unroll
    for (int l = 0; l < LA; ++l) {
      if (tid + l * 256 < BM * 4) {
        As[buf][a_kq[l] * 4 + 0][a_r[l]] = ra[l].x;
        As[buf][a_kq[l] * 4 + 1][a_r[l]] = ra[l].y;
        As[buf][a_kq[l] * 4 + 2][a_r[l]] = ra[l].z;
        As[buf][a_kq[l] * 4 + 3][a_r[l]] = ra[l].w;
      }
    }
#pragma unroll
    for (int l = 0; l < LW; ++l) {
      if (tid + l * 256 < BN * 4) {
        Ws[buf][w_kq[l] * 4 + 0][w_r[l]] = rw[l].x;
        Ws[buf][w_kq[l] * 4 + 1][w_r[l]] = rw[l].y;
        Ws[buf][w_kq[l] * 4 + 2][w_r[l]] = rw[l].z;
        Ws[buf][w_kq[l] * 4 + 3][w_r[l]] = rw[l].w;
      }
    }
  };

  float acc[4][4];
#pragma unroll
  for (int i = 0; i < 4; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) acc[i][j] = 0.0f;

  const int nk = K / BK;
  gload(0);
  sstore(0);
  __syncthreads();
  for (int kt = 0; kt < nk; ++kt) {
    const int buf = kt & 1;
    if (kt + 1 < nk) gload((kt + 1) * BK);
#pragma unroll
    for (int k = 0; k < BK; ++k) {
      const float4 a = *reinterpret_cast<const float4*>(&As[buf][k][ty * 4]);
      const float4 w = *reinterpret_cast<const float4*>(&Ws[buf][k][tx * 4]);
      const float av[4] = {a.x, a.y, a.z, a.w};
      const float wv[4] = {w.x, w.y, w.z, w.w};
#pragma unroll
      for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(av[i], wv[j], acc[i][j]);
    }
    if (kt + 1 < nk) {
      sstore(buf ^ 1);  // other buffer: its readers finished at the barrier of iteration kt-1
      __syncthreads();
    }
  }

#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int m = m0 + ty * 4 + i;
    if (m < M && n0 + tx * 4 < N) epilogue_store4<TO>(ep, m, n0 + tx * 4, N, acc[i]);
  }
}

template <typename TA, typename TO>
static cudaError_t launch_typed(const GemmProblem& p, cudaStream_t st) {
  const TA* A = reinterpret_cast<const TA*>(p.A);
  const TA* W = reinterpret_cast<const TA*>(p.W);
  if (p.N <= 32) {
    dim3 grid((p.N + 31) / 32, (p.M + 127) / 128);
    launch_k(gemm_simt_kernel<TA, TO, 128, 32>, grid, 256, 0, st, A, p.a_map, W, p.M, p.N, p.K, p.ep);
  } else if (p.M <= 32) {
    dim3 grid((p.N + 127) / 128, (p.M + 31) / 32);
    launch_k(gemm_simt_kernel<TA, TO, 32, 128>, grid, 256, 0, st, A, p.a_map, W, p.M, p.N, p.K, p.ep);
  } else {
    dim3 grid((p.N + 63) / 64, (p.M + 63) / 64);
    launch_k(gemm_simt_kernel<TA, TO, 64, 64>, grid, 256, 0, st, A, p.a_map, W, p.M, p.N, p.K, p.ep);
  }
  return cudaGetLastError();
}

cudaError_t gemm_simt_launch(const GemmProblem& p, int in_type, int out_type, cudaStream_t st) {
  if (p.M <= 0 || p.N <= 0) return cudaSuccess;
  if (p.K % 16 != 0) return cudaErrorInvalidValue;
  if (in_type == 0 && out_type == 0) return launch_typed<float, float>(p, st);
  if (in_type == 1 && out_type == 0) return launch_typed<bf16, float>(p, st);
  if (in_type == 1 && out_type == 1) return launch_typed<bf16, bf16>(p, st);
  return cudaErrorInvalidValue;
}

}  // namespace ptts
