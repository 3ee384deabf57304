// Flow-head cluster kernel (bf16 mode): one Euler step of the LSD sampler for up to 64 streams in ONE launch.
//
//   u = input_proj(x)                                              mlp.py:203
//   6 x ResBlock: u += gate * mlp2(SiLU(mlp0(LN(u) (1 + scale) + shift)))   mlp.py:86-111
//   x' = x + alpha * final(LN_noaffine(u) (1 + scale) + shift)      mlp.py:114-131, flow_lm.py:32-40
//   (optionally) commit x' as the stream's previous latent          flow_lm.py:138-139, tts_model.py:767-768
//
// The AdaLN modulation rows `mod` (shift | scale | gate per block, shift | scale for the final layer) come from
// the wide weight-streaming GEMM that runs before this kernel; everything after it used to be 21 dependent
// launches of ~7 us each for a few MFLOP.  Here a thread-block cluster of 8 CTAs owns 64 streams: CTA c holds
// features [64c, 64c+64) of the 512-wide residual `u` in registers (mma.sync accumulator layout), computes its
// 64 output features of every 512x512 GEMM with mma.sync.m16n8k16 (streams = M, features = N), and pushes the
// bf16 operand slice of the next GEMM into all 8 CTAs' shared memory through DSMEM.  LayerNorm statistics are
// exchanged as per-warp (sum, M2) pairs and combined with the parallel-variance formula.  The only
// synchronisation is barrier.cluster (3 per ResBlock).  Weights are read straight from L2 into registers in a
// K order that matches a 16-byte shared-memory load of the activations (the contraction index may be permuted
// as long as both operands agree), so no weight staging exists; the CTA's 64 KB slice of every matrix is
// prefetched into L2 before the kernel waits for its producer (programmatic dependent launch).
#include "kernels.cuh"
#include "tc_ptx.cuh"

namespace ptts {

namespace {

constexpr int HC_CL = 8;          // CTAs per cluster
constexpr int HC_FS = 64;         // features per CTA
constexpr int HC_F = HC_CL * HC_FS;
constexpr int HC_ROWS = 64;       // streams per cluster
constexpr int HC_XP = 544;        // bf16 per operand row (1088 B: 16 B loads of 8 rows x 64 B hit distinct banks)
constexpr int HC_LP = 72;         // bf16 per staging row
constexpr int HC_THREADS = 256;
constexpr int HC_PD = 4;          // weight prefetch distance in 32-wide K chunks
constexpr int HC_KC = HC_F / 32;  // K chunks per GEMM

struct HeadSmem {
  bf16 X[2][HC_ROWS][HC_XP];
  bf16 loc[HC_ROWS][HC_LP];
  float2 part[2 * HC_CL][HC_ROWS];
  float2 stat[HC_ROWS];
};

__device__ __forceinline__ void cluster_sync_all() {
  cluster_arrive_release();
  cluster_wait_acquire();
}
__device__ __forceinline__ uint32_t cluster_rank() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
  return r;
}
__device__ __forceinline__ void dsmem_st_v4(uint32_t addr, uint4 v) {
  asm volatile("st.shared::cluster.v4.b32 [%0], {%1,%2,%3,%4};" ::"r"(addr), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
}
__device__ __forceinline__ void dsmem_st_f2(uint32_t addr, float x, float y) {
  asm volatile("st.shared::cluster.v2.f32 [%0], {%1,%2};" ::"r"(addr), "f"(x), "f"(y) : "memory");
}
__device__ __forceinline__ uint4 ldg_nc_v4(const void* p) {
  uint4 r;
  asm volatile("ld.global.nc.v4.b32 {%0,%1,%2,%3}, [%4];" : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p));
  return r;
}
__device__ __forceinline__ void prefetch_l2(const void* p) { asm volatile("prefetch.global.L2 [%0];" ::"l"(p)); }
__device__ __forceinline__ void mma16816(float (&c)[4], uint32_t a0, uint32_t a1, uint32_t a2, uint32_t a3, uint32_t b0, uint32_t b1) {
  asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
               : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
               : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
}
__device__ __forceinline__ uint32_t pack2(float lo, float hi) {
  const __nv_bfloat162 b = __floats2bfloat162_rn(lo, hi);
  return *reinterpret_cast<const uint32_t*>(&b);
}

// Weight prefetch registers of one GEMM: HC_PD chunks x NT feature tiles x 16 bytes per lane.
template <int NT>
__device__ __forceinline__ void w_prefetch(uint4 (&wq)[HC_PD][4], const bf16* wrow /* W + (tile base + g) * K + 8 t */) {
#pragma unroll
  for (int p = 0; p < HC_PD; ++p)
#pragma unroll
    for (int j = 0; j < NT; ++j) wq[p][j] = ldg_nc_v4(wrow + (size_t)j * 8 * HC_F + 32 * p);
}

// acc[j] (16 streams x 8 features per tile j) = sum_k X[stream][k] * W[feature][k] over K = 512.
// Lane (g, t) reads the 16 bytes at k = 32 kc + 8 t .. + 7 of rows g / g + 8 (activations) and of feature row g of
// each tile (weights): two k16 MMAs per chunk whose logical k columns {2t, 2t+1, 2t+8, 2t+9} map to the physical
// columns {8t..8t+3} and {8t+4..8t+7} in both operands.
template <int NT>
__device__ __forceinline__ void head_mma(float (&acc)[4][4], uint4 (&wq)[HC_PD][4], const bf16* wrow, const bf16* xa, const bf16* xb) {
#pragma unroll
  for (int j = 0; j < NT; ++j) acc[j][0] = acc[j][1] = acc[j][2] = acc[j][3] = 0.0f;
#pragma unroll
  for (int kc = 0; kc < HC_KC; ++kc) {
    uint4 w[NT];
#pragma unroll
    for (int j = 0; j < NT; ++j) w[j] = wq[kc % HC_PD][j];
    if (kc + HC_PD < HC_KC) {
#pragma unroll
      for (int j = 0; j < NT; ++j) wq[kc % HC_PD][j] = ldg_nc_v4(wrow + (size_t)j * 8 * HC_F + 32 * (kc + HC_PD));
    }
    const uint4 a = *reinterpret_cast<const uint4*>(xa + 32 * kc);
    const uint4 b = *reinterpret_cast<const uint4*>(xb + 32 * kc);
#pragma unroll
    for (int j = 0; j < NT; ++j) {
      mma16816(acc[j], a.x, b.x, a.y, b.y, w[j].x, w[j].y);
      mma16816(acc[j], a.z, b.z, a.w, b.w, w[j].z, w[j].w);
    }
  }
}

__global__ void __launch_bounds__(HC_THREADS, 1)
flow_head_cluster_kernel(const __grid_constant__ FlowHeadParams P) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  HeadSmem& S = *reinterpret_cast<HeadSmem*>(smem_raw);
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int g = lane >> 2, t = lane & 3;
  const int mw = warp & 3, nh = warp >> 2;
  const uint32_t c = cluster_rank();
  const int q = blockIdx.x / HC_CL;                 // cluster index = 64-stream group
  const int row0 = q * HC_ROWS;
  const int sa = 16 * mw + g, sb = sa + 8;          // this thread's streams (rows of the cluster)
  const bool va = row0 + sa < P.n, vb = row0 + sb < P.n;
  const int fl0 = 32 * nh + 2 * t;                  // local feature of tile j, element e: fl0 + 8 j + e
  const int f0 = HC_FS * (int)c + fl0;              // global feature
  const int depth = P.depth;

  pdl_launch_dependents();
  // ---- before the producer kernel is known to be complete: only weights are touched
  for (int m = 0; m < 2 * depth; ++m) {             // this CTA's 64 KB slice of every 512 x 512 matrix -> L2
    const bf16* w = ((m & 1) ? P.blk[m >> 1].w2 : P.blk[m >> 1].w0) + (size_t)HC_FS * c * HC_F;
    for (int i = tid; i < HC_FS * HC_F * 2 / 128; i += HC_THREADS) prefetch_l2(reinterpret_cast<const char*>(w) + (size_t)i * 128);
  }
  if (c == 0)
    for (int i = tid; i < P.ldim * HC_F * 2 / 128; i += HC_THREADS) prefetch_l2(reinterpret_cast<const char*>(P.w_final) + (size_t)i * 128);
  uint4 wq[HC_PD][4];
  const bf16* wrow = P.blk[0].w0 + (size_t)(HC_FS * c + 32 * nh + g) * HC_F + 8 * t;
  w_prefetch<4>(wq, wrow);
  {  // input_proj weights of this CTA's features -> shared (fp32, pitch 33) in the second operand buffer
    float* win = reinterpret_cast<float*>(&S.X[1][0][0]) + HC_ROWS * 33;
    for (int i = tid; i < HC_FS * P.ldim; i += HC_THREADS) {
      const int f = i / P.ldim, k = i - f * P.ldim;
      win[f * 33 + k] = P.w_in[(size_t)(HC_FS * c + f) * P.ldim + k];
    }
  }
  cluster_sync_all();   // every CTA of the cluster is resident before any DSMEM traffic
  pdl_wait();

  // ---- u = input_proj(x) in fp32 (K = ldim = 32)
  float u[2][4][2];     // [stream a/b][tile j][element e]
  {
    float* xin = reinterpret_cast<float*>(&S.X[1][0][0]);
    float* win = xin + HC_ROWS * 33;
    for (int i = tid; i < HC_ROWS * P.ldim; i += HC_THREADS) {
      const int s = i / P.ldim, k = i - s * P.ldim;
      xin[s * 33 + k] = (row0 + s < P.n) ? P.x_in[(size_t)(row0 + s) * P.ldim + k] : 0.0f;
    }
    __syncthreads();
#pragma unroll
    for (int j = 0; j < 4; ++j)
#pragma unroll
      for (int e = 0; e < 2; ++e) {
        const float b = P.b_in[f0 + 8 * j + e];
        u[0][j][e] = b; u[1][j][e] = b;
      }
    for (int k = 0; k < P.ldim; ++k) {
      const float xa = xin[sa * 33 + k], xb = xin[sb * 33 + k];
#pragma unroll
      for (int j = 0; j < 4; ++j)
#pragma unroll
        for (int e = 0; e < 2; ++e) {
          const float w = win[(fl0 + 8 * j + e) * 33 + k];
          u[0][j][e] = fmaf(xa, w, u[0][j][e]);
          u[1][j][e] = fmaf(xb, w, u[1][j][e]);
        }
    }
    // the region is reused as a bf16 operand buffer only after the next two cluster barriers
  }

  const uint32_t part_base = smem_u32(&S.part[0][0]);
  const float* mod_a = P.mod + (size_t)(row0 + sa) * P.mod_ld;
  const float* mod_b = P.mod + (size_t)(row0 + sb) * P.mod_ld;
  const bf16* xa_rd[2] = {&S.X[0][sa][8 * t], &S.X[1][sa][8 * t]};
  const bf16* xb_rd[2] = {&S.X[0][sb][8 * t], &S.X[1][sb][8 * t]};

  // LayerNorm statistics of u over the 512 features of each stream: per-warp (sum, M2) over 32 features pushed to
  // every CTA, cluster barrier, parallel-variance combine -> S.stat[stream] = (mean, sqrt(var + eps)).
  auto exchange_stats = [&]() {
    float s_a = 0.0f, s_b = 0.0f;
#pragma unroll
    for (int j = 0; j < 4; ++j) { s_a += u[0][j][0] + u[0][j][1]; s_b += u[1][j][0] + u[1][j][1]; }
    s_a += __shfl_xor_sync(0xffffffffu, s_a, 1); s_a += __shfl_xor_sync(0xffffffffu, s_a, 2);
    s_b += __shfl_xor_sync(0xffffffffu, s_b, 1); s_b += __shfl_xor_sync(0xffffffffu, s_b, 2);
    const float m_a = s_a * (1.0f / 32.0f), m_b = s_b * (1.0f / 32.0f);
    float q_a = 0.0f, q_b = 0.0f;
#pragma unroll
    for (int j = 0; j < 4; ++j)
#pragma unroll
      for (int e = 0; e < 2; ++e) {
        const float da = u[0][j][e] - m_a, db = u[1][j][e] - m_b;
        q_a = fmaf(da, da, q_a); q_b = fmaf(db, db, q_b);
      }
    q_a += __shfl_xor_sync(0xffffffffu, q_a, 1); q_a += __shfl_xor_sync(0xffffffffu, q_a, 2);
    q_b += __shfl_xor_sync(0xffffffffu, q_b, 1); q_b += __shfl_xor_sync(0xffffffffu, q_b, 2);
    const uint32_t pa = part_base + (uint32_t)(((2 * c + nh) * HC_ROWS + sa) * sizeof(float2));
    const uint32_t pb = part_base + (uint32_t)(((2 * c + nh) * HC_ROWS + sb) * sizeof(float2));
#pragma unroll
    for (int d = 0; d < 2; ++d) {                   // lane t serves destination CTAs 2t and 2t + 1
      const uint32_t dst = 2 * t + d;
      dsmem_st_f2(dsmem_addr(pa, dst), s_a, q_a);
      dsmem_st_f2(dsmem_addr(pb, dst), s_b, q_b);
    }
    cluster_sync_all();
    {
      const int s = tid >> 2, qq = tid & 3;         // 4 lanes per stream, 4 partials each
      float2 pv[4];
      float tot = 0.0f;
#pragma unroll
      for (int i = 0; i < 4; ++i) { pv[i] = S.part[4 * qq + i][s]; tot += pv[i].x; }
      tot += __shfl_xor_sync(0xffffffffu, tot, 1); tot += __shfl_xor_sync(0xffffffffu, tot, 2);
      const float mu = tot * (1.0f / HC_F);
      float m2 = 0.0f;
#pragma unroll
      for (int i = 0; i < 4; ++i) { const float dm = pv[i].x * (1.0f / 32.0f) - mu; m2 += pv[i].y + 32.0f * dm * dm; }
      m2 += __shfl_xor_sync(0xffffffffu, m2, 1); m2 += __shfl_xor_sync(0xffffffffu, m2, 2);
      if (qq == 0) S.stat[s] = make_float2(mu, sqrtf(m2 * (1.0f / HC_F) + 1e-6f));
    }
    __syncthreads();
  };

  // bf16 slice (64 streams x this CTA's 64 features) staged in S.loc -> column block 64c of X[buf] in all 8 CTAs
  auto broadcast = [&](int buf) {
    __syncthreads();
    const int r = tid >> 2, ch = (tid & 3) * 2;
    const uint4 v0 = *reinterpret_cast<const uint4*>(&S.loc[r][8 * ch]);
    const uint4 v1 = *reinterpret_cast<const uint4*>(&S.loc[r][8 * ch + 8]);
    const uint32_t dst = smem_u32(&S.X[buf][r][HC_FS * c + 8 * ch]);
#pragma unroll
    for (int d = 0; d < HC_CL; ++d) {
      const uint32_t a = dsmem_addr(dst, (c + d) & (HC_CL - 1));
      dsmem_st_v4(a, v0);
      dsmem_st_v4(a + 16, v1);
    }
    cluster_sync_all();
  };

  for (int blk = 0; blk < depth; ++blk) {
    const FlowHeadBlock& B = P.blk[blk];
    // per-stream modulation rows and per-feature vectors of this block (in flight across the statistics exchange)
    float2 sh[2][4], sc[2][4], gt[2][4], lw[4], lb[4], b0[4], b2[4];
    const int mo = blk * 3 * HC_F;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int f = f0 + 8 * j;
      const float2 z = make_float2(0.0f, 0.0f);
      sh[0][j] = va ? load2(mod_a + mo + f) : z;             sh[1][j] = vb ? load2(mod_b + mo + f) : z;
      sc[0][j] = va ? load2(mod_a + mo + HC_F + f) : z;      sc[1][j] = vb ? load2(mod_b + mo + HC_F + f) : z;
      gt[0][j] = va ? load2(mod_a + mo + 2 * HC_F + f) : z;  gt[1][j] = vb ? load2(mod_b + mo + 2 * HC_F + f) : z;
      lw[j] = load2(B.lnw + f); lb[j] = load2(B.lnb + f); b0[j] = load2(B.b0 + f); b2[j] = load2(B.b2 + f);
    }
    exchange_stats();
    {
      const float2 st_a = S.stat[sa], st_b = S.stat[sb];
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const float a0 = ((u[0][j][0] - st_a.x) / st_a.y * lw[j].x + lb[j].x) * (1.0f + sc[0][j].x) + sh[0][j].x;
        const float a1 = ((u[0][j][1] - st_a.x) / st_a.y * lw[j].y + lb[j].y) * (1.0f + sc[0][j].y) + sh[0][j].y;
        const float c0 = ((u[1][j][0] - st_b.x) / st_b.y * lw[j].x + lb[j].x) * (1.0f + sc[1][j].x) + sh[1][j].x;
        const float c1 = ((u[1][j][1] - st_b.x) / st_b.y * lw[j].y + lb[j].y) * (1.0f + sc[1][j].y) + sh[1][j].y;
        *reinterpret_cast<uint32_t*>(&S.loc[sa][fl0 + 8 * j]) = pack2(a0, a1);
        *reinterpret_cast<uint32_t*>(&S.loc[sb][fl0 + 8 * j]) = pack2(c0, c1);
      }
    }
    broadcast(0);
    float acc[4][4];
    head_mma<4>(acc, wq, wrow, xa_rd[0], xb_rd[0]);                       // mlp0
    wrow = B.w2 + (size_t)(HC_FS * c + 32 * nh + g) * HC_F + 8 * t;
    w_prefetch<4>(wq, wrow);
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      *reinterpret_cast<uint32_t*>(&S.loc[sa][fl0 + 8 * j]) = pack2(silu_call(acc[j][0] + b0[j].x), silu_call(acc[j][1] + b0[j].y));
      *reinterpret_cast<uint32_t*>(&S.loc[sb][fl0 + 8 * j]) = pack2(silu_call(acc[j][2] + b0[j].x), silu_call(acc[j][3] + b0[j].y));
    }
    broadcast(1);
    head_mma<4>(acc, wq, wrow, xa_rd[1], xb_rd[1]);                       // mlp2
    if (blk + 1 < depth) {
      wrow = P.blk[blk + 1].w0 + (size_t)(HC_FS * c + 32 * nh + g) * HC_F + 8 * t;
      w_prefetch<4>(wq, wrow);
    } else if (c == 0) {
      wrow = P.w_final + (size_t)(16 * nh + g) * HC_F + 8 * t;
      w_prefetch<2>(wq, wrow);
    }
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      u[0][j][0] = fmaf(gt[0][j].x, acc[j][0] + b2[j].x, u[0][j][0]);
      u[0][j][1] = fmaf(gt[0][j].y, acc[j][1] + b2[j].y, u[0][j][1]);
      u[1][j][0] = fmaf(gt[1][j].x, acc[j][2] + b2[j].x, u[1][j][0]);
      u[1][j][1] = fmaf(gt[1][j].y, acc[j][3] + b2[j].y, u[1][j][1]);
    }
  }

  // ---- final layer: LN without affine, modulate, Linear 512 -> ldim, Euler update
  {
    float2 sh[2][4], sc[2][4];
    const int mo = depth * 3 * HC_F;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int f = f0 + 8 * j;
      const float2 z = make_float2(0.0f, 0.0f);
      sh[0][j] = va ? load2(mod_a + mo + f) : z;         sh[1][j] = vb ? load2(mod_b + mo + f) : z;
      sc[0][j] = va ? load2(mod_a + mo + HC_F + f) : z;  sc[1][j] = vb ? load2(mod_b + mo + HC_F + f) : z;
    }
    exchange_stats();
    const float2 st_a = S.stat[sa], st_b = S.stat[sb];
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const float a0 = (u[0][j][0] - st_a.x) / st_a.y * (1.0f + sc[0][j].x) + sh[0][j].x;
      const float a1 = (u[0][j][1] - st_a.x) / st_a.y * (1.0f + sc[0][j].y) + sh[0][j].y;
      const float c0 = (u[1][j][0] - st_b.x) / st_b.y * (1.0f + sc[1][j].x) + sh[1][j].x;
      const float c1 = (u[1][j][1] - st_b.x) / st_b.y * (1.0f + sc[1][j].y) + sh[1][j].y;
      *reinterpret_cast<uint32_t*>(&S.loc[sa][fl0 + 8 * j]) = pack2(a0, a1);
      *reinterpret_cast<uint32_t*>(&S.loc[sb][fl0 + 8 * j]) = pack2(c0, c1);
    }
    broadcast(0);
    if (c != 0) return;   // no DSMEM access after the last cluster barrier: the other CTAs may retire
    float acc[4][4];
    head_mma<2>(acc, wq, wrow, xa_rd[0], xb_rd[0]);
#pragma unroll
    for (int j = 0; j < 2; ++j) {
      const int o = 16 * nh + 8 * j + 2 * t;
      const float2 bf = load2(P.b_final + o);
#pragma unroll
      for (int r = 0; r < 2; ++r) {
        const int s = row0 + (r ? sb : sa);
        if (s >= P.n) continue;
        const float2 xi = load2(P.x_in + (size_t)s * P.ldim + o);
        const float2 v = make_float2(fmaf(P.alpha, acc[j][2 * r] + bf.x, xi.x), fmaf(P.alpha, acc[j][2 * r + 1] + bf.y, xi.y));
        store2(P.x_out + (size_t)s * P.ldim + o, v);
        if (P.prev_latent) {
          store2(P.prev_latent + (size_t)P.slot_ids[s] * P.ldim + o, v);
          if (P.commit_out) store2(P.commit_out + (size_t)s * P.ldim + o, v);
        }
      }
    }
  }
}

}  // namespace

bool flow_head_cluster_supported(int flow_dim, int ldim, int depth) {
  return flow_dim == HC_F && ldim == 32 && depth >= 1 && depth <= PTTS_HEAD_MAX_DEPTH;
}

cudaError_t launch_flow_head_cluster(const FlowHeadParams& p, cudaStream_t st) {
  static DeviceOnce attr_done;
  const size_t smem = sizeof(HeadSmem);
  if (attr_done.first()) {
    cudaError_t e = cudaFuncSetAttribute(flow_head_cluster_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
  }
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = dim3(HC_CL * ((p.n + HC_ROWS - 1) / HC_ROWS), 1, 1);
  cfg.blockDim = dim3(HC_THREADS, 1, 1);
  cfg.dynamicSmemBytes = smem;
  cfg.stream = st;
  cudaLaunchAttribute attr[2];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = HC_CL;
  attr[0].val.clusterDim.y = 1;
  attr[0].val.clusterDim.z = 1;
  attr[1].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[1].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = pdl_enabled() ? 2 : 1;
  return cudaLaunchKernelEx(&cfg, flow_head_cluster_kernel, p);
}

}  // namespace ptts
