// Flow-head cluster kernel, tcgen05 form (bf16 mode, default): one Euler step of the LSD sampler for up to 64
// streams per cluster in ONE launch — input_proj, the AdaLN ResBlocks, the final layer, the latent commit
// (mlp.py:86-131,188-215; flow_lm.py:32-40).  Same decomposition as flow_head.cu (8-CTA cluster, CTA c owns
// features [64c, 64c+64) of the 512-wide residual), different machinery:
//
//  * every 512x512 GEMM slice is 8 k-blocks of tcgen05.mma (M = 128 TMEM lanes of which 64 are stream rows,
//    N = 64 features, K = 64 per block) on SWIZZLE_128B operands; the accumulator lives in TMEM;
//  * weights stream through an 8-stage ring of 8 KB TMA tiles that runs ahead across GEMM boundaries (the
//    producer starts before the kernel waits for its predecessor: weights depend on nothing);
//  * the activation operand of the next GEMM is exchanged without any cluster barrier: each CTA writes its
//    64 x 64 bf16 slice in the canonical swizzled layout — which is exactly k-block c of the operand — and one
//    thread pushes it to the 8 CTAs with cp.async.bulk (shared::cta -> shared::cluster) completing on the
//    destination's mbarrier; LayerNorm partials travel the same way with st.async;
//  * 4 epilogue warps own one stream row each (TMEM lane = stream), 32 features per thread, so LayerNorm
//    statistics, AdaLN modulation, SiLU, gating and the residual are register-local.
//
// Buffer reuse needs no extra synchronisation: a CTA can only reach the next LayerNorm exchange after every peer
// has sent its partials, i.e. after every peer has finished reading both operand buffers of the previous block.
#include <cstdio>
#include <cstdlib>

#include "gemm.cuh"
#include "kernels.cuh"
#include "tc_ptx.cuh"

namespace ptts {

PTTS_TL_ENABLE_FN(head_tc)

namespace {

constexpr int HT_CL = 8;             // CTAs per cluster
constexpr int HT_FS = 64;            // features per CTA = K per k-block
constexpr int HT_F = HT_CL * HT_FS;  // 512
constexpr int HT_ROWS = 64;          // streams per cluster
constexpr int HT_THREADS = 192;      // warps 0,1,4,5: rows / epilogue (TMEM lane quarters 0,1); warp 2: TMA; warp 3: MMA
constexpr int HT_STAGES = 8;
constexpr uint32_t HT_BLK = HT_ROWS * 128;          // 8 KB: one k-block of the activation operand / one weight stage
constexpr uint32_t HT_XBYTES = HT_CL * HT_BLK;      // 64 KB per operand buffer
constexpr uint32_t HT_TMEM_COLS = 64;
constexpr int HT_XP = 36;            // fp32 pitch of the input_proj staging rows

// shared memory carve (offsets from a 1024-byte aligned base)
constexpr uint32_t OFF_X = 0;                                  // [2][8][64 rows][128 B]
constexpr uint32_t OFF_W = OFF_X + 2 * HT_XBYTES;              // [8 stages][64 rows][128 B]
constexpr uint32_t OFF_LOC = OFF_W + HT_STAGES * HT_BLK;       // [2][64 rows][128 B]
constexpr uint32_t OFF_PART = OFF_LOC + 2 * HT_BLK;            // [16 partials][64 streams] float2
constexpr uint32_t OFF_VEC = OFF_PART + 2 * HT_CL * HT_ROWS * 8;   // [depth][4][64] fp32: lnw, lnb, b0, b2
constexpr uint32_t OFF_BAR = OFF_VEC + PTTS_HEAD_MAX_DEPTH * 4 * HT_FS * 4;
constexpr uint32_t HT_SMEM = OFF_BAR + 256 + 1024;

struct Bars {
  uint64_t wfull[HT_STAGES], wempty[HT_STAGES], xfull[2], accfull, stat;
  uint32_t tmem_base;
};

__device__ __forceinline__ uint32_t cluster_rank() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
  return r;
}
// bounded wait: a protocol error traps (CUDA error at the next sync) instead of hanging the GPU
__device__ __forceinline__ void wait_bar(uint32_t bar, uint32_t parity) {
  for (uint32_t spin = 0; spin < (1u << 22); ++spin)
    if (mbar_test_a(bar, parity)) return;
  __trap();
}
__device__ __forceinline__ void bulk_smem_to_cluster(uint32_t dst_cluster, uint32_t src_cta, uint32_t bytes, uint32_t bar_cluster) {
  asm volatile("cp.async.bulk.shared::cluster.shared::cta.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
               ::"r"(dst_cluster), "r"(src_cta), "r"(bytes), "r"(bar_cluster) : "memory");
}
// one TMA tile (64 k x 64 rows, SWIZZLE_128B) from global memory into the same shared-memory offset of every CTA
// in `mask`, completing on the mbarrier at the same offset in each of them
__device__ __forceinline__ void tma_load_2d_mcast(uint32_t dst, const CUtensorMap* map, uint32_t bar, int c0, int c1, uint16_t mask) {
  asm volatile(
      "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes.multicast::cluster [%0], [%1, {%3, %4}], [%2], %5;"
      ::"r"(dst), "l"(map), "r"(bar), "r"(c0), "r"(c1), "h"(mask) : "memory");
}
__device__ __forceinline__ void st_async_f2(uint32_t dst_cluster, float x, float y, uint32_t bar_cluster) {
  asm volatile("st.async.shared::cluster.mbarrier::complete_tx::bytes.v2.f32 [%0], {%1, %2}, [%3];"
               ::"r"(dst_cluster), "f"(x), "f"(y), "r"(bar_cluster) : "memory");
}
__device__ __forceinline__ void prefetch_l2(const void* p) { asm volatile("prefetch.global.L2 [%0];" ::"l"(p)); }
__device__ __forceinline__ void epi_bar() { asm volatile("bar.sync 1, 128;" ::: "memory"); }
__device__ __forceinline__ uint32_t pack2(float lo, float hi) {
  const __nv_bfloat162 b = __floats2bfloat162_rn(lo, hi);
  return *reinterpret_cast<const uint32_t*>(&b);
}

// developer trace (PTTS_HEAD_TRACE=1, eager mode): globaltimer stamps of CTA 0's first row thread
__device__ unsigned long long g_head_trace[128];
__device__ __forceinline__ unsigned long long gtimer() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  return t;
}
#define HT_TRACE(i) do { if (P.trace && c == 0 && blockIdx.x < HT_CL && tid == 0) g_head_trace[i] = gtimer(); } while (0)

__global__ void __launch_bounds__(HT_THREADS, 1)
flow_head_tc_kernel(const __grid_constant__ FlowHeadTcParams P) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* sm = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  const uint32_t sm_a = smem_u32(sm);
  Bars& bars = *reinterpret_cast<Bars*>(sm + OFF_BAR);
  float* vec = reinterpret_cast<float*>(sm + OFF_VEC);
  float2* part = reinterpret_cast<float2*>(sm + OFF_PART);
  const FlowHeadParams& H = P.h;

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const uint32_t c = cluster_rank();
  const int row0 = (blockIdx.x / HT_CL) * HT_ROWS;
  const int depth = H.depth;
  const int n_gemm = 2 * depth;                     // plus the final layer on CTA 0
  const uint32_t wfull_a = smem_u32(&bars.wfull[0]), wempty_a = smem_u32(&bars.wempty[0]);
  const uint32_t xfull_a = smem_u32(&bars.xfull[0]), accfull_a = smem_u32(&bars.accfull), stat_a = smem_u32(&bars.stat);

  pdl_launch_dependents();
  const unsigned tl = tl_begin(6);
  HT_TRACE(0);
  // ---- prologue: nothing here depends on the previous kernel
  for (int m = 0; m < n_gemm; ++m) {                // this CTA's 64 KB slice of every 512 x 512 matrix -> L2
    const bf16* w = ((m & 1) ? H.blk[m >> 1].w2 : H.blk[m >> 1].w0) + (size_t)HT_FS * c * HT_F;
    for (int i = tid; i < HT_FS * HT_F * 2 / 128; i += HT_THREADS) prefetch_l2(reinterpret_cast<const char*>(w) + (size_t)i * 128);
  }
  if (c == 0)
    for (int i = tid; i < H.ldim * HT_F * 2 / 128; i += HT_THREADS) prefetch_l2(reinterpret_cast<const char*>(H.w_final) + (size_t)i * 128);
  if (tid == 0) {
    for (int s = 0; s < HT_STAGES; ++s) { mbar_init(&bars.wfull[s], 1); mbar_init(&bars.wempty[s], 1); }
    mbar_init(&bars.xfull[0], 1); mbar_init(&bars.xfull[1], 1);
    mbar_init(&bars.accfull, 1); mbar_init(&bars.stat, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 3) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&bars.tmem_base)), "n"(HT_TMEM_COLS) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  for (int i = tid; i < depth * 4 * HT_FS; i += HT_THREADS) {   // per-feature vectors of this CTA's features
    const int b = i / (4 * HT_FS), v = (i / HT_FS) & 3, f = i & (HT_FS - 1);
    const FlowHeadBlock& B = H.blk[b];
    const float* src = v == 0 ? B.lnw : v == 1 ? B.lnb : v == 2 ? B.b0 : B.b2;
    vec[i] = src[HT_FS * c + f];
  }
  float* xin = reinterpret_cast<float*>(sm + OFF_X + HT_XBYTES);   // input_proj staging lives in operand buffer 1
  float* win = xin + HT_ROWS * HT_XP;
  for (int i = tid; i < HT_FS * H.ldim; i += HT_THREADS) {
    const int f = i / H.ldim, k = i - f * H.ldim;
    win[f * HT_XP + k] = H.w_in[(size_t)(HT_FS * c + f) * H.ldim + k];
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  cluster_arrive_release();   // every CTA of the cluster is resident with initialised barriers before any
  cluster_wait_acquire();     // remote traffic
  const uint32_t tmem_base = bars.tmem_base;
  HT_TRACE(1);

  if (warp == 2) {
    // ===== weight producer: all k-blocks of all GEMMs in program order through the ring =====
    if (lane == 0) {
      for (int m = 0; m < n_gemm + 1; ++m) asm volatile("prefetch.tensormap [%0];" ::"l"(&P.maps[m]) : "memory");
      uint32_t s = 0, ph = 0;
      const int total = n_gemm + (c == 0 ? 1 : 0);
      for (int m = 0; m < total; ++m) {
        const bool fin = m == n_gemm;
        const uint32_t bytes = fin ? (uint32_t)H.ldim * 128u : HT_BLK;
        const int wrow = fin ? 0 : HT_FS * (int)c;
        for (int kb = 0; kb < HT_CL; ++kb) {
          wait_bar(wempty_a + 8u * s, ph ^ 1u);
          mbar_expect_tx_a(wfull_a + 8u * s, bytes);
          tma_load_2d_a(sm_a + OFF_W + s * HT_BLK, &P.maps[m], wfull_a + 8u * s, kb * HT_FS, wrow);
          if (++s == HT_STAGES) { s = 0; ph ^= 1u; }
        }
      }
    }
  } else if (warp == 3) {
    // ===== MMA issuer =====
    if (lane == 0) {
      // instruction descriptor: D = f32, A = B = bf16, both K-major, N >> 3 at bit 17, M >> 4 at bit 24
      const uint32_t idesc64 = (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(64 >> 3) << 17) | ((uint32_t)(128 >> 4) << 24);
      const uint32_t idescL = (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(H.ldim >> 3) << 17) | ((uint32_t)(128 >> 4) << 24);
      uint32_t s = 0, ph = 0, xph[2] = {0, 0};
      const int total = n_gemm + 1;                  // the last operand exchange (final layer) is awaited by every CTA
      for (int m = 0; m < total; ++m) {
        const int buf = m & 1;
        mbar_expect_tx_a(xfull_a + 8u * buf, HT_XBYTES);
        wait_bar(xfull_a + 8u * buf, xph[buf]);
        xph[buf] ^= 1u;
        if (m == n_gemm && c != 0) break;            // only CTA 0 computes the final layer
        tc_fence_after();
        const uint32_t idesc = m == n_gemm ? idescL : idesc64;
        const uint64_t adesc0 = make_sw128_desc(sm_a + OFF_X + buf * HT_XBYTES);
        for (int kb = 0; kb < HT_CL; ++kb) {
          wait_bar(wfull_a + 8u * s, ph);
          tc_fence_after();
          const uint64_t adesc = adesc0 + (uint64_t)((kb * HT_BLK) >> 4);
          const uint64_t bdesc = make_sw128_desc(sm_a + OFF_W + s * HT_BLK);
#pragma unroll
          for (int k = 0; k < HT_FS / 16; ++k)
            tc_mma_bf16(tmem_base, adesc + (uint64_t)(k * 2), bdesc + (uint64_t)(k * 2), idesc, (kb | k) != 0);
          tc_commit_a(wempty_a + 8u * s);
          if (++s == HT_STAGES) { s = 0; ph ^= 1u; }
        }
        tc_commit_a(accfull_a);
      }
    }
  } else {
    // ===== row warps: one stream per thread, 32 of this CTA's 64 features =====
    const int qd = warp & 3, ch = warp >> 2;         // TMEM lane quarter (0, 1), feature half
    const int et = (qd * 32 + lane) + 64 * ch;       // 0..127
    const int s = 32 * qd + lane;                    // stream row inside the cluster
    const bool valid = row0 + s < H.n;
    const int fl0 = 32 * ch;                         // first local feature of this thread
    const int f0 = HT_FS * (int)c + fl0;             // first global feature
    const uint32_t taddr = tmem_base + ((uint32_t)(qd * 32) << 16) + (uint32_t)fl0;
    uint32_t accph = 0, statph = 0;
    pdl_wait();                                      // x_in / mod come from the previous kernels
    tl_mark(tl, 2);
    HT_TRACE(2);

    float u[32];
    {  // u = input_proj(x): fp32, K = ldim
      for (int i = et; i < HT_ROWS * H.ldim; i += 128) {
        const int r = i / H.ldim, k = i - r * H.ldim;
        xin[r * HT_XP + k] = (row0 + r < H.n) ? H.x_in[(size_t)(row0 + r) * H.ldim + k] : 0.0f;
      }
      epi_bar();
      float x[32];
#pragma unroll
      for (int k = 0; k < 32; k += 4) {
        const float4 v = *reinterpret_cast<const float4*>(xin + s * HT_XP + k);
        x[k] = v.x; x[k + 1] = v.y; x[k + 2] = v.z; x[k + 3] = v.w;
      }
#pragma unroll
      for (int i = 0; i < 32; ++i) {
        float a = H.b_in[f0 + i];
#pragma unroll
        for (int k = 0; k < 32; k += 4) {
          const float4 w = *reinterpret_cast<const float4*>(win + (fl0 + i) * HT_XP + k);
          a = fmaf(x[k], w.x, a); a = fmaf(x[k + 1], w.y, a); a = fmaf(x[k + 2], w.z, a); a = fmaf(x[k + 3], w.w, a);
        }
        u[i] = a;
      }
    }
    const float* mod_row = H.mod + (size_t)(row0 + s) * H.mod_ld;
    HT_TRACE(3);

    // LayerNorm statistics over the 512 features of this thread's stream: (sum, M2) of its 32 features to every
    // CTA (st.async completes on the destination's `stat` barrier), then the parallel-variance combine.
    auto ln_stats = [&](float& mu, float& den) {
      float sum = 0.0f;
#pragma unroll
      for (int i = 0; i < 32; ++i) sum += u[i];
      const float m = sum * (1.0f / 32.0f);
      float m2 = 0.0f;
#pragma unroll
      for (int i = 0; i < 32; ++i) { const float d = u[i] - m; m2 = fmaf(d, d, m2); }
      if (et == 0) mbar_expect_tx_a(stat_a, 2 * HT_CL * HT_ROWS * 8);
      const uint32_t slot = sm_a + OFF_PART + (uint32_t)(((2 * c + ch) * HT_ROWS + s) * 8);
#pragma unroll
      for (int d = 0; d < HT_CL; ++d) {
        const uint32_t dst = (c + d) & (HT_CL - 1);
        st_async_f2(dsmem_addr(slot, dst), sum, m2, dsmem_addr(stat_a, dst));
      }
      wait_bar(stat_a, statph);
      statph ^= 1u;
      float2 pv[2 * HT_CL];
      float tot = 0.0f;
#pragma unroll
      for (int i = 0; i < 2 * HT_CL; ++i) { pv[i] = part[i * HT_ROWS + s]; tot += pv[i].x; }
      mu = tot * (1.0f / HT_F);
      float q = 0.0f;
#pragma unroll
      for (int i = 0; i < 2 * HT_CL; ++i) { const float dm = pv[i].x * (1.0f / 32.0f) - mu; q += pv[i].y + 32.0f * dm * dm; }
      den = sqrtf(q * (1.0f / HT_F) + 1e-6f);
    };
    // g[32] (this thread's row, features fl0..fl0+31) -> bf16, swizzled k-block layout in loc[buf] -> k-block c of
    // operand buffer `buf` in all 8 CTAs
    auto publish = [&](const float* gv, int buf) {
      if (P.xchg) {
        // through L2: the slice goes to a global scratch block as plain rows, then ONE multicast TMA load brings
        // it (swizzled by the tensor map) into k-block c of the operand buffer of all 8 CTAs.  DSMEM pushes of
        // the same 8 x 8 KB measured ~4 us per exchange; an L2 round trip with multicast is ~1.5 us.
        const int blk_row = (((blockIdx.x / HT_CL) * 2 + buf) * HT_CL + (int)c) * HT_ROWS;
        bf16* dstg = P.xchg + (size_t)(blk_row + s) * HT_FS + fl0;
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          uint4 v;
          v.x = pack2(gv[8 * j], gv[8 * j + 1]); v.y = pack2(gv[8 * j + 2], gv[8 * j + 3]);
          v.z = pack2(gv[8 * j + 4], gv[8 * j + 5]); v.w = pack2(gv[8 * j + 6], gv[8 * j + 7]);
          *reinterpret_cast<uint4*>(dstg + 8 * j) = v;
        }
        fence_proxy_async();
        tc_fence_before();
        epi_bar();
        if (et == 0)
          tma_load_2d_mcast(sm_a + OFF_X + buf * HT_XBYTES + c * HT_BLK, &P.xmap, xfull_a + 8u * buf, 0, blk_row, (uint16_t)0xFF);
        return;
      }
      uint8_t* loc = sm + OFF_LOC + buf * HT_BLK + s * 128;
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        uint4 v;
        v.x = pack2(gv[8 * j], gv[8 * j + 1]); v.y = pack2(gv[8 * j + 2], gv[8 * j + 3]);
        v.z = pack2(gv[8 * j + 4], gv[8 * j + 5]); v.w = pack2(gv[8 * j + 6], gv[8 * j + 7]);
        *reinterpret_cast<uint4*>(loc + (((4 * ch + j) ^ (s & 7)) << 4)) = v;
      }
      fence_proxy_async();
      tc_fence_before();
      epi_bar();
      if (et == 0) {
        const uint32_t src = sm_a + OFF_LOC + buf * HT_BLK;
        const uint32_t dst = sm_a + OFF_X + buf * HT_XBYTES + c * HT_BLK;
#pragma unroll
        for (int d = 0; d < HT_CL; ++d) {
          const uint32_t r = (c + d) & (HT_CL - 1);
          bulk_smem_to_cluster(dsmem_addr(dst, r), src, HT_BLK, dsmem_addr(xfull_a + 8u * buf, r));
        }
      }
    };

    for (int blk = 0; blk < depth; ++blk) {
      const float* vb = vec + blk * 4 * HT_FS + fl0;   // lnw | lnb | b0 | b2, 64 floats apart
      const int mo = blk * 3 * HT_F + f0;
      float sh[32], sc[32];
#pragma unroll
      for (int i = 0; i < 32; i += 4) {
        const float4 a = valid ? ld4f(mod_row + mo + i) : make_float4(0, 0, 0, 0);
        const float4 b = valid ? ld4f(mod_row + mo + HT_F + i) : make_float4(0, 0, 0, 0);
        sh[i] = a.x; sh[i + 1] = a.y; sh[i + 2] = a.z; sh[i + 3] = a.w;
        sc[i] = b.x; sc[i + 1] = b.y; sc[i + 2] = b.z; sc[i + 3] = b.w;
      }
      float mu, den;
      ln_stats(mu, den);
      HT_TRACE(4 + blk * 5);
      float gv[32];
#pragma unroll
      for (int i = 0; i < 32; ++i)
        gv[i] = ((u[i] - mu) / den * vb[i] + vb[HT_FS + i]) * (1.0f + sc[i]) + sh[i];
      publish(gv, 0);
      HT_TRACE(5 + blk * 5);
      float gt[32];                                    // gate row: in flight across both GEMMs
#pragma unroll
      for (int i = 0; i < 32; i += 4) {
        const float4 a = valid ? ld4f(mod_row + mo + 2 * HT_F + i) : make_float4(0, 0, 0, 0);
        gt[i] = a.x; gt[i + 1] = a.y; gt[i + 2] = a.z; gt[i + 3] = a.w;
      }
      wait_bar(accfull_a, accph); accph ^= 1u;         // mlp.0
      HT_TRACE(6 + blk * 5);
      tc_fence_after();
      tmem_ld<32>(taddr, gv);
#pragma unroll
      for (int i = 0; i < 32; ++i) gv[i] = silu_call(gv[i] + vb[2 * HT_FS + i]);
      publish(gv, 1);
      HT_TRACE(7 + blk * 5);
      wait_bar(accfull_a, accph); accph ^= 1u;         // mlp.2
      HT_TRACE(8 + blk * 5);
      tc_fence_after();
      tmem_ld<32>(taddr, gv);
#pragma unroll
      for (int i = 0; i < 32; ++i) u[i] = fmaf(gt[i], gv[i] + vb[3 * HT_FS + i], u[i]);
      tc_fence_before();
    }
    {  // final layer: LN without affine, modulate, Linear 512 -> ldim on CTA 0, Euler update, commit
      const int mo = depth * 3 * HT_F + f0;
      float sh[32], sc[32];
#pragma unroll
      for (int i = 0; i < 32; i += 4) {
        const float4 a = valid ? ld4f(mod_row + mo + i) : make_float4(0, 0, 0, 0);
        const float4 b = valid ? ld4f(mod_row + mo + HT_F + i) : make_float4(0, 0, 0, 0);
        sh[i] = a.x; sh[i + 1] = a.y; sh[i + 2] = a.z; sh[i + 3] = a.w;
        sc[i] = b.x; sc[i + 1] = b.y; sc[i + 2] = b.z; sc[i + 3] = b.w;
      }
      float mu, den;
      ln_stats(mu, den);
      float gv[32];
#pragma unroll
      for (int i = 0; i < 32; ++i) gv[i] = (u[i] - mu) / den * (1.0f + sc[i]) + sh[i];
      publish(gv, 0);
      if (c == 0 && ch == 0) {
        wait_bar(accfull_a, accph);
        tc_fence_after();
        tmem_ld<32>(taddr, gv);
        const int sg = row0 + s;
        if (valid) {
          const float* xi = H.x_in + (size_t)sg * H.ldim;
          float* xo = H.x_out + (size_t)sg * H.ldim;
          float* pl = H.prev_latent ? H.prev_latent + (size_t)H.slot_ids[sg] * H.ldim : nullptr;
          float* co = (H.prev_latent && H.commit_out) ? H.commit_out + (size_t)sg * H.ldim : nullptr;
#pragma unroll
          for (int i = 0; i < 32; i += 4) {
            const float4 x4 = ld4f(xi + i), b4 = ld4f(H.b_final + i);
            float4 o;
            o.x = fmaf(H.alpha, gv[i] + b4.x, x4.x); o.y = fmaf(H.alpha, gv[i + 1] + b4.y, x4.y);
            o.z = fmaf(H.alpha, gv[i + 2] + b4.z, x4.z); o.w = fmaf(H.alpha, gv[i + 3] + b4.w, x4.w);
            *reinterpret_cast<float4*>(xo + i) = o;
            if (pl) *reinterpret_cast<float4*>(pl + i) = o;
            if (co) *reinterpret_cast<float4*>(co + i) = o;
          }
        }
        tc_fence_before();
      }
    }
  }
  // ---- every CTA has received the last operand exchange (MMA warp waited for it) before anyone retires: no
  // bulk copy is in flight into or out of a CTA that has left
  HT_TRACE(60);
  __syncthreads();
  cluster_arrive_release();
  cluster_wait_acquire();
  HT_TRACE(61);
  tl_mark(tl, 3);
  if (warp == 3) {
    tc_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "n"(HT_TMEM_COLS) : "memory");
  }
}

}  // namespace

bool flow_head_tc_prepare(FlowHeadTcParams* p) {
  const FlowHeadParams& h = p->h;
  if (p->xchg) {
    const long long rows = (long long)((h.n + HT_ROWS - 1) / HT_ROWS) * 2 * HT_CL * HT_ROWS;
    if (!tc_make_map(&p->xmap, p->xchg, rows, HT_FS, HT_FS, HT_ROWS)) return false;
  }
  for (int m = 0; m < 2 * h.depth; ++m) {
    const bf16* w = (m & 1) ? h.blk[m >> 1].w2 : h.blk[m >> 1].w0;
    if (!tc_make_map(&p->maps[m], w, HT_F, HT_F, HT_F, HT_FS)) return false;
  }
  return tc_make_map(&p->maps[2 * h.depth], h.w_final, h.ldim, HT_F, HT_F, h.ldim);
}

cudaError_t launch_flow_head_tc(const FlowHeadTcParams& p, cudaStream_t st) {
  static DeviceOnce attr_done;
  if (attr_done.first()) {
    cudaError_t e = cudaFuncSetAttribute(flow_head_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)HT_SMEM);
    if (e != cudaSuccess) return e;
  }
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = dim3(HT_CL * ((p.h.n + HT_ROWS - 1) / HT_ROWS), 1, 1);
  cfg.blockDim = dim3(HT_THREADS, 1, 1);
  cfg.dynamicSmemBytes = HT_SMEM;
  cfg.stream = st;
  cudaLaunchAttribute attr[2];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = HT_CL;
  attr[0].val.clusterDim.y = 1;
  attr[0].val.clusterDim.z = 1;
  attr[1].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[1].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = pdl_enabled() ? 2 : 1;
  static int trace = -1, calls = 0;
  if (trace < 0) { const char* e = getenv("PTTS_HEAD_TRACE"); trace = (e && e[0] == '1') ? 1 : 0; }
  if (!trace) return cudaLaunchKernelEx(&cfg, flow_head_tc_kernel, p);
  FlowHeadTcParams q = p;
  q.trace = 1;
  cudaError_t e = cudaLaunchKernelEx(&cfg, flow_head_tc_kernel, q);
  cudaStreamCaptureStatus cs = cudaStreamCaptureStatusNone;
  cudaStreamIsCapturing(st, &cs);
  if (e == cudaSuccess && cs == cudaStreamCaptureStatusNone && ++calls == 8) {
    unsigned long long t[128];
    cudaStreamSynchronize(st);
    cudaMemcpyFromSymbol(t, g_head_trace, sizeof t);
    fprintf(stderr, "[head trace] ns since kernel start:");
    for (int i = 1; i < 62; ++i) if (t[i]) fprintf(stderr, " %d:%lld", i, (long long)(t[i] - t[0]));
    fprintf(stderr, "\n");
  }
  return e;
}

}  // namespace ptts
