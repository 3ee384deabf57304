// Flow-head kernel, stream-split form (bf16 mode, default): one Euler step of the LSD sampler (mlp.py:86-131,188-215;
// flow_lm.py:32-40) with NO cross-CTA exchange.  Each CTA owns 8 streams (SS_N) and computes the whole 512-wide head for
// them; every CTA streams all the head weights (12 matrices of 512 x 512 bf16 + the final 32 x 512) through a TMA ring
// — they are L2-resident after the first CTA touched them.
//
// Why mma.sync and not tcgen05 here (measured, profiles/r02_mma_sync_probe.txt, r02_l2_ingest_probe2.txt and the first
// tcgen05 form of this kernel): with 8-16 streams per CTA the GEMMs are [512 x 512] x [512 x 8..16].  tcgen05.mma reads
// its shared-memory A operand (128 rows x 16 k = 4 KB) at 32 B/clk, i.e. 128 cycles per instruction whatever N is, so
// with the weights as A one SM consumes weights at 63 GB/s (8.3 us per matrix, measured); with the weights as B
// (N = 256) it is 64 B/clk = 4.2 us per matrix and the accumulator wastes 112 of 128 lanes.  mma.sync.m16n8k16 runs at
// 0.5 instructions/clk/SM with its A fragment fed by ldmatrix.x4 (128 B/clk of weights), the same rate at which two
// producer warps can fill the ring from L2 (one issuing thread sustains one bulk-tensor copy per 0.24 us whatever its
// size; 2 x 32 KB boxes reach the SM's ~124 B/clk ingest port).  Measured: ~4 us per 512 x 512 matrix — the TMA fill
// (512 KB written) and the ldmatrix reads (512 KB) share the shared-memory port; the whole step 65 us after its wait
// against 97 us for the 8-CTA feature-split cluster kernel (flow_head_tc.cu) it replaces.
//
//   warps 0..7    compute: warp w owns features 32 w .. 32 w + 31 of each 256-row half (4 m16 tiles) for all streams
//                 (one n8 tile): 16 fp32 accumulators per thread; the residual u, the AdaLN shift / scale / gate and the
//                 LayerNorm all live in the same fragment layout, so bias, SiLU, gate and residual are register-local.
//   warps 8, 9    producers (stage i is issued by warp 8 + i % 2): 5 stages of 32 KB = [256 weight rows][64 k]
//                 SWIZZLE_128B, k-block-major order so that a k-block's activation fragments are loaded once for both
//                 halves.  They start before the programmatic-dependent-launch wait (weights depend on nothing) and
//                 run ahead across GEMM boundaries.  The fp32 input_proj weight rides through the first two stages.
//   LayerNorm     statistics over the 512 features of a stream: quad shuffles + one named barrier (CTA-local).
//   operand       each warp writes its features of the next GEMM's activation operand (bf16, [stream][k] rows of
//                 128 B, SWIZZLE_128B-style XOR so that ldmatrix is conflict-free) and one named barrier hands it over.
#include <cstdio>
#include <cstdlib>

#include "gemm.cuh"
#include "kernels.cuh"
#include "tc_ptx.cuh"

namespace ptts {

PTTS_TL_ENABLE_FN(head_ss)

namespace {

constexpr int SS_NT = 1;                 // n8 stream tiles per CTA (1 or 2)
constexpr int SS_N = 8 * SS_NT;          // streams per CTA
constexpr int SS_F = 512;                // head width
constexpr int SS_KB = SS_F / 64;         // k-blocks of 64 (128 B of bf16)
constexpr int SS_CW = 8;                 // compute warps
constexpr int SS_PROD = 2;               // producer warps
constexpr int SS_THREADS = (SS_CW + SS_PROD) * 32;
constexpr int SS_STAGES = 5;
constexpr int SS_STAGE_ROWS = 256;
constexpr uint32_t SS_STAGE_BYTES = SS_STAGE_ROWS * 128;   // 32 KB
constexpr int SS_WIN_STAGES = 2;         // fp32 input_proj weight [512][32] = 64 KB = two stages
constexpr uint32_t SS_KB_BYTES = SS_N * 128;            // one k-block of the activation operand (SS_N rows x 128 B)
constexpr uint32_t SS_OP_BYTES = SS_KB * SS_KB_BYTES;
constexpr int SS_LD = 32;                // latent width (checked by the launcher)

constexpr uint32_t SO_OP = 0;                                        // [8 kb][SS_N rows][128 B]
constexpr uint32_t SO_RING = SO_OP + SS_OP_BYTES;                    // [5][256 rows][128 B]
constexpr uint32_t SO_X = SO_RING + SS_STAGES * SS_STAGE_BYTES;      // [16][32] fp32
constexpr uint32_t SO_STAT = SO_X + SS_N * SS_LD * 4;                // [2][8 warps][SS_N streams] float2
constexpr uint32_t SO_BAR = SO_STAT + 2 * SS_CW * SS_N * 8;
constexpr uint32_t SO_VEC = SO_BAR + 256;                            // [depth][lnw | lnb | b0 | b2][512], then b_in [512], b_final [32] (fp32)
__host__ __device__ constexpr uint32_t ss_vec_floats(int depth) { return (uint32_t)(depth * 4 * SS_F + SS_F + SS_LD); }
__host__ __device__ constexpr uint32_t ss_smem_bytes(int depth) { return SO_VEC + ss_vec_floats(depth) * 4 + 1024; }

struct SsBars {
  uint64_t wfull[SS_STAGES], wempty[SS_STAGES];
};

__device__ __forceinline__ void wait_bar(uint32_t bar, uint32_t parity) {   // bounded: a protocol error traps
  for (uint32_t spin = 0; spin < (1u << 22); ++spin)
    if (mbar_test_a(bar, parity)) return;
  __trap();
}
__device__ __forceinline__ void cw_bar() { asm volatile("bar.sync 1, 256;" ::: "memory"); }   // the 8 compute warps
__device__ __forceinline__ void bulk_g2s(uint32_t dst, const void* src, uint32_t bytes, uint32_t bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
               ::"r"(dst), "l"(src), "r"(bytes), "r"(bar) : "memory");
}
__device__ __forceinline__ void ldsm4(uint32_t* r, uint32_t addr) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0,%1,%2,%3}, [%4];"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]) : "r"(addr));
}
__device__ __forceinline__ void mma16816(float* d, const uint32_t* a, uint32_t b0, uint32_t b1) {
  asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
               : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3]) : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}

__device__ unsigned long long g_ss_trace[64];
#define SS_TRACE(i) do { if (P.trace && blockIdx.x == 0 && tid == 0) g_ss_trace[i] = tl_now(); } while (0)

__global__ void __launch_bounds__(SS_THREADS, 1)
flow_head_ss_kernel(const __grid_constant__ FlowHeadTcParams P) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* sm = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  const uint32_t sm_a = smem_u32(sm);
  SsBars& bars = *reinterpret_cast<SsBars*>(sm + SO_BAR);
  const FlowHeadParams& H = P.h;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int row0 = blockIdx.x * SS_N;
  const int depth = H.depth, n_gemm = 2 * depth;
  const uint32_t wfull_a = smem_u32(&bars.wfull[0]), wempty_a = smem_u32(&bars.wempty[0]);

  pdl_launch_dependents();
  const unsigned tl = tl_begin(6);
  SS_TRACE(0);
  if (tid == 0) {
    for (int s = 0; s < SS_STAGES; ++s) { mbar_init(&bars.wfull[s], 1); mbar_init(&bars.wempty[s], SS_CW); }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();

  if (warp >= SS_CW) {
    // ===== producers: input_proj weight (fp32, 2 stages), then the 32 KB stages of every GEMM in consumption order
    // (k-block major, the two 256-row halves inside), then the 8 k-blocks of the final 32 x 512 matrix =====
    if (lane == 0) {
      const int me = warp - SS_CW;
      if (me == 0)
        for (int m = 0; m < n_gemm + 1; ++m) asm volatile("prefetch.tensormap [%0];" ::"l"(&P.maps[m]) : "memory");
      uint32_t i = 0;                                   // global stage sequence number
      for (; i < SS_WIN_STAGES; ++i)
        if ((int)(i % SS_PROD) == me) {
          mbar_expect_tx_a(wfull_a + 8u * i, SS_STAGE_BYTES);
          bulk_g2s(sm_a + SO_RING + i * SS_STAGE_BYTES, reinterpret_cast<const char*>(H.w_in) + (size_t)i * SS_STAGE_BYTES,
                   SS_STAGE_BYTES, wfull_a + 8u * i);
        }
      for (int m = 0; m < n_gemm; ++m)
        for (int kb = 0; kb < SS_KB; ++kb)
          for (int hb = 0; hb < 2; ++hb, ++i) {
            if ((int)(i % SS_PROD) != me) continue;
            const uint32_t s = i % SS_STAGES, ph = (i / SS_STAGES) & 1u;
            wait_bar(wempty_a + 8u * s, ph ^ 1u);
            mbar_expect_tx_a(wfull_a + 8u * s, SS_STAGE_BYTES);
            tma_load_2d_a(sm_a + SO_RING + s * SS_STAGE_BYTES, &P.maps[m], wfull_a + 8u * s, kb * 64, hb * SS_STAGE_ROWS);
          }
      if ((int)(i % SS_PROD) == me) {                   // final 32 x 512 matrix: ONE stage of 8 k-block boxes of 4 KB
        const uint32_t s = i % SS_STAGES, ph = (i / SS_STAGES) & 1u;
        wait_bar(wempty_a + 8u * s, ph ^ 1u);
        mbar_expect_tx_a(wfull_a + 8u * s, (uint32_t)SS_KB * SS_LD * 128u);
        for (int kb = 0; kb < SS_KB; ++kb)
          tma_load_2d_a(sm_a + SO_RING + s * SS_STAGE_BYTES + kb * (SS_LD * 128), &P.maps[n_gemm], wfull_a + 8u * s, kb * 64, 0);
      }
    }
    return;
  }

  // ===== compute warps =====
  // accumulator / residual layout (mma.sync m16n8 C fragment): value [mt][nt][e] is
  //   feature f(mt, e) = 256 (mt >> 1) + 32 warp + 16 (mt & 1) + lane / 4 + 8 (e >> 1)
  //   stream  j(nt, e) = 8 nt + 2 (lane % 4) + (e & 1)
  const int g4 = lane >> 2, t4 = lane & 3;
  float* xs = reinterpret_cast<float*>(sm + SO_X);
  float2* stat = reinterpret_cast<float2*>(sm + SO_STAT);
  int stat_buf = 0;
  uint32_t seq = SS_WIN_STAGES;                         // stage sequence number (same on all compute warps)

  auto feat = [&](int mt, int e) { return 256 * (mt >> 1) + 32 * warp + 16 * (mt & 1) + g4 + 8 * (e >> 1); };
  // A fragment address pieces: lane l addresses row (l & 7) + 8 ((l >> 3) & 1) of the 16-row tile, 16-byte chunk (l >> 4)
  const uint32_t a_row = (uint32_t)(32 * warp + (lane & 7) + 8 * ((lane >> 3) & 1));
  const uint32_t a_chunk = (uint32_t)(lane >> 4);
  // B fragment (activations [stream][k]): two n8 tiles: one ldmatrix.x4 per k16 step, lane l addresses stream row
  // (l & 7) + 8 (l >> 4), chunk (l >> 3) & 1; one n8 tile: one ldmatrix.x4 per two k16 steps, row l & 7, chunk l >> 3
  const uint32_t b_row = SS_NT == 2 ? (uint32_t)((lane & 7) + 8 * (lane >> 4)) : (uint32_t)(lane & 7);
  const uint32_t b_chunk = SS_NT == 2 ? (uint32_t)((lane >> 3) & 1) : (uint32_t)(lane >> 3);
  const uint32_t b_base = sm_a + SO_OP + (b_row >> 3) * 1024u + (b_row & 7u) * 128u;

  // one GEMM: acc[mt][nt][e] = sum_k W[f][k] * op[j][k] over the stages seq .. (rows `tiles` m16 tiles per half)
  float acc[4][SS_NT][4];
  auto gemm = [&]() {
#pragma unroll
    for (int mt = 0; mt < 4; ++mt)
#pragma unroll
      for (int nt = 0; nt < SS_NT; ++nt)
#pragma unroll
        for (int e = 0; e < 4; ++e) acc[mt][nt][e] = 0.0f;
#pragma unroll 1
    for (int kb = 0; kb < SS_KB; ++kb) {
      uint32_t bfr[4][2 * SS_NT];                       // [k16 step][b0, b1 per n8 tile]
      if (SS_NT == 2) {
#pragma unroll
        for (int kk = 0; kk < 4; ++kk)
          ldsm4(&bfr[kk][0], b_base + (uint32_t)kb * SS_KB_BYTES + ((((uint32_t)(2 * kk) + b_chunk) ^ (b_row & 7u)) << 4));
      } else {
#pragma unroll
        for (int k2 = 0; k2 < 2; ++k2) {
          uint32_t r4[4];
          ldsm4(r4, b_base + (uint32_t)kb * SS_KB_BYTES + ((((uint32_t)(4 * k2) + b_chunk) ^ (b_row & 7u)) << 4));
          bfr[2 * k2][0] = r4[0]; bfr[2 * k2][1] = r4[1]; bfr[2 * k2 + 1][0] = r4[2]; bfr[2 * k2 + 1][1] = r4[3];
        }
      }
#pragma unroll
      for (int hb = 0; hb < 2; ++hb) {
        const uint32_t s = seq % SS_STAGES, ph = (seq / SS_STAGES) & 1u;
        wait_bar(wfull_a + 8u * s, ph);
        const uint32_t st = sm_a + SO_RING + s * SS_STAGE_BYTES;
#pragma unroll
        for (int kk = 0; kk < 4; ++kk)
#pragma unroll
          for (int mi = 0; mi < 2; ++mi) {
            const uint32_t r = a_row + 16u * mi;
            uint32_t a[4];
            ldsm4(a, st + r * 128u + ((((uint32_t)(2 * kk) + a_chunk) ^ (r & 7u)) << 4));
#pragma unroll
            for (int nt = 0; nt < SS_NT; ++nt) mma16816(acc[2 * hb + mi][nt], a, bfr[kk][2 * nt], bfr[kk][2 * nt + 1]);
          }
        __syncwarp();
        if (lane == 0) mbar_arrive_a(wempty_a + 8u * s);
        ++seq;
      }
    }
  };
  // operand element (stream j, feature f) as bf16 for the next GEMM
  auto op_store = [&](int f, int j, float v) {
    const uint32_t off = SO_OP + (uint32_t)(f >> 6) * SS_KB_BYTES + (uint32_t)(j >> 3) * 1024u + (uint32_t)(j & 7) * 128u +
                         ((((uint32_t)(f & 63) >> 3) ^ (uint32_t)(j & 7)) << 4) + (uint32_t)(f & 7) * 2u;
    *reinterpret_cast<bf16*>(sm + off) = __float2bfloat16_rn(v);
  };

  // per-feature vectors of every block (weights: independent of the previous kernel) -> shared memory
  float* vec = reinterpret_cast<float*>(sm + SO_VEC);
  float* v_bin = vec + depth * 4 * SS_F;
  float* v_bfin = v_bin + SS_F;
  for (int i = tid; i < depth * 4 * (SS_F / 4); i += SS_CW * 32) {
    const int b = i / (4 * (SS_F / 4)), v = (i / (SS_F / 4)) & 3, c = i % (SS_F / 4);
    const FlowHeadBlock& B = H.blk[b];
    const float* src = v == 0 ? B.lnw : v == 1 ? B.lnb : v == 2 ? B.b0 : B.b2;
    reinterpret_cast<float4*>(vec)[i] = __ldg(reinterpret_cast<const float4*>(src) + c);
  }
  for (int i = tid; i < SS_F; i += SS_CW * 32) v_bin[i] = __ldg(H.b_in + i);
  if (tid < SS_LD) v_bfin[tid] = __ldg(H.b_final + tid);
  pdl_wait();                                           // x_in / mod come from the previous kernels
  tl_mark(tl, 2);
  SS_TRACE(1);
  for (int i = tid; i < SS_N * SS_LD; i += SS_CW * 32) {
    const int r = i / SS_LD;
    xs[i] = (row0 + r < H.n) ? H.x_in[(size_t)row0 * SS_LD + i] : 0.0f;
  }
  bool valid[SS_NT][2];
  const float* mod_row[SS_NT][2];
#pragma unroll
  for (int nt = 0; nt < SS_NT; ++nt)
#pragma unroll
    for (int b = 0; b < 2; ++b) {
      const int j = 8 * nt + 2 * t4 + b;
      valid[nt][b] = row0 + j < H.n;
      mod_row[nt][b] = H.mod + (size_t)(valid[nt][b] ? row0 + j : 0) * H.mod_ld;
    }
  // modulation vector at column offset `off` in this thread's fragment layout
  auto load_mod = [&](float (*dst)[SS_NT][4], int off) {
#pragma unroll
    for (int mt = 0; mt < 4; ++mt)
#pragma unroll
      for (int nt = 0; nt < SS_NT; ++nt)
#pragma unroll
        for (int e = 0; e < 4; ++e)
          dst[mt][nt][e] = valid[nt][e & 1] ? __ldg(mod_row[nt][e & 1] + off + feat(mt, e)) : 0.0f;
  };
  float sh[4][SS_NT][4], sc[4][SS_NT][4];
  load_mod(sh, 0);
  load_mod(sc, SS_F);
  cw_bar();

  // ---- u = input_proj(x): fp32 weight rows wait in ring stages 0, 1 (feature f at byte f * 128)
  float u[4][SS_NT][4];
#pragma unroll
  for (int mt = 0; mt < 4; ++mt)
#pragma unroll
    for (int e = 0; e < 4; e += 2) {
      const float b = v_bin[feat(mt, e)];
#pragma unroll
      for (int nt = 0; nt < SS_NT; ++nt) { u[mt][nt][e] = b; u[mt][nt][e + 1] = b; }
    }
  for (int t = 0; t < SS_WIN_STAGES; ++t) wait_bar(wfull_a + 8u * t, 0);
#pragma unroll 2
  for (int c = 0; c < 8; ++c) {
    const int kk = (c + g4) & 7;                        // rotated k order: the 8 feature rows of a quad column hit 8 bank groups
    float4 xv[SS_NT][2];
#pragma unroll
    for (int nt = 0; nt < SS_NT; ++nt)
#pragma unroll
      for (int b = 0; b < 2; ++b) xv[nt][b] = *reinterpret_cast<const float4*>(xs + (8 * nt + 2 * t4 + b) * SS_LD + 4 * kk);
#pragma unroll
    for (int mt = 0; mt < 4; ++mt)
#pragma unroll
      for (int h2 = 0; h2 < 2; ++h2) {
        const float4 w = *reinterpret_cast<const float4*>(sm + SO_RING + feat(mt, 2 * h2) * 128 + kk * 16);
#pragma unroll
        for (int nt = 0; nt < SS_NT; ++nt)
#pragma unroll
          for (int b = 0; b < 2; ++b) {
            float& d = u[mt][nt][2 * h2 + b];
            d = fmaf(w.x, xv[nt][b].x, fmaf(w.y, xv[nt][b].y, fmaf(w.z, xv[nt][b].z, fmaf(w.w, xv[nt][b].w, d))));
          }
      }
  }
  __syncwarp();
  if (lane == 0)
    for (int t = 0; t < SS_WIN_STAGES; ++t) mbar_arrive_a(wempty_a + 8u * t);
  SS_TRACE(2);

  // LayerNorm statistics of the 16 streams over 512 features: per-thread partials over its 8 features of 4 streams,
  // shuffles over the 8 lanes that share a quad column, the 8 warps combined through shared memory.
  auto ln_stats = [&](float (*mu)[2], float (*rs)[2]) {
    float2* sb = stat + stat_buf * (SS_CW * SS_N);
#pragma unroll
    for (int nt = 0; nt < SS_NT; ++nt)
#pragma unroll
      for (int b = 0; b < 2; ++b) {
        float s1 = 0.0f, s2 = 0.0f;
#pragma unroll
        for (int mt = 0; mt < 4; ++mt)
#pragma unroll
          for (int h2 = 0; h2 < 2; ++h2) { const float v = u[mt][nt][2 * h2 + b]; s1 += v; s2 = fmaf(v, v, s2); }
#pragma unroll
        for (int o = 4; o < 32; o <<= 1) { s1 += __shfl_xor_sync(0xffffffffu, s1, o); s2 += __shfl_xor_sync(0xffffffffu, s2, o); }
        if (g4 == 0) sb[warp * SS_N + 8 * nt + 2 * t4 + b] = make_float2(s1, s2);
      }
    cw_bar();
#pragma unroll
    for (int nt = 0; nt < SS_NT; ++nt)
#pragma unroll
      for (int b = 0; b < 2; ++b) {
        float s1 = 0.0f, s2 = 0.0f;
#pragma unroll
        for (int w = 0; w < SS_CW; ++w) { const float2 p = sb[w * SS_N + 8 * nt + 2 * t4 + b]; s1 += p.x; s2 += p.y; }
        const float mean = s1 * (1.0f / SS_F);
        mu[nt][b] = mean;
        rs[nt][b] = rsqrtf(fmaxf(s2 * (1.0f / SS_F) - mean * mean, 0.0f) + 1e-6f);
      }
    stat_buf ^= 1;
  };

  for (int blk = 0; blk < depth; ++blk) {
    const float* vb = vec + blk * 4 * SS_F;
    float mu[SS_NT][2], rs[SS_NT][2];
    ln_stats(mu, rs);
    // g = LN(u) * (1 + scale) + shift -> operand of mlp.0.  (The previous GEMM's reads of the operand buffer ended
    // before the barrier inside ln_stats.)
#pragma unroll
    for (int mt = 0; mt < 4; ++mt)
#pragma unroll
      for (int h2 = 0; h2 < 2; ++h2) {
        const int f = feat(mt, 2 * h2);
        const float lw = vb[f], lb = vb[SS_F + f];
#pragma unroll
        for (int nt = 0; nt < SS_NT; ++nt)
#pragma unroll
          for (int b = 0; b < 2; ++b) {
            const int e = 2 * h2 + b;
            op_store(f, 8 * nt + 2 * t4 + b, fmaf(fmaf((u[mt][nt][e] - mu[nt][b]) * rs[nt][b], lw, lb), 1.0f + sc[mt][nt][e], sh[mt][nt][e]));
          }
      }
    // gate of this block and shift | scale of the next one: in flight across both GEMMs
    float gt[4][SS_NT][4];
    const int mo = blk * 3 * SS_F;
    load_mod(gt, mo + 2 * SS_F);
    load_mod(sh, mo + 3 * SS_F);
    load_mod(sc, mo + 4 * SS_F);
    cw_bar();                                           // operand complete
    SS_TRACE(3 + blk * 3);
    gemm();                                        // mlp.0
    cw_bar();                                           // every warp has read the operand: it may be overwritten
#pragma unroll
    for (int mt = 0; mt < 4; ++mt)
#pragma unroll
      for (int h2 = 0; h2 < 2; ++h2) {
        const int f = feat(mt, 2 * h2);
        const float b0 = vb[2 * SS_F + f];
#pragma unroll
        for (int nt = 0; nt < SS_NT; ++nt)
#pragma unroll
          for (int b = 0; b < 2; ++b) op_store(f, 8 * nt + 2 * t4 + b, silu_call(acc[mt][nt][2 * h2 + b] + b0));
      }
    cw_bar();
    SS_TRACE(4 + blk * 3);
    gemm();                                        // mlp.2
#pragma unroll
    for (int mt = 0; mt < 4; ++mt)
#pragma unroll
      for (int h2 = 0; h2 < 2; ++h2) {
        const float b2 = vb[3 * SS_F + feat(mt, 2 * h2)];
#pragma unroll
        for (int nt = 0; nt < SS_NT; ++nt)
#pragma unroll
          for (int b = 0; b < 2; ++b) {
            const int e = 2 * h2 + b;
            u[mt][nt][e] = fmaf(gt[mt][nt][e], acc[mt][nt][e] + b2, u[mt][nt][e]);
          }
      }
    SS_TRACE(5 + blk * 3);
  }
  {  // final layer: LN without affine, modulate, Linear 512 -> ldim, Euler update, commit
    float mu[SS_NT][2], rs[SS_NT][2];
    ln_stats(mu, rs);
#pragma unroll
    for (int mt = 0; mt < 4; ++mt)
#pragma unroll
      for (int nt = 0; nt < SS_NT; ++nt)
#pragma unroll
        for (int e = 0; e < 4; ++e)
          op_store(feat(mt, e), 8 * nt + 2 * t4 + (e & 1), fmaf((u[mt][nt][e] - mu[nt][e & 1]) * rs[nt][e & 1], 1.0f + sc[mt][nt][e], sh[mt][nt][e]));
    cw_bar();
    // Linear 512 -> 32: the whole matrix is one stage (8 k-block boxes of 32 rows); warp w takes k-block w, the 8
    // partial [32 x streams] tiles are summed through shared memory (a drained ring stage)
    const uint32_t s = seq % SS_STAGES, ph = (seq / SS_STAGES) & 1u;
    wait_bar(wfull_a + 8u * s, ph);
    float fa[2][SS_NT][4] = {};
    {
      const uint32_t st = sm_a + SO_RING + s * SS_STAGE_BYTES + (uint32_t)warp * (SS_LD * 128);
      uint32_t bfr[4][2 * SS_NT];
      if (SS_NT == 2) {
#pragma unroll
        for (int kk = 0; kk < 4; ++kk)
          ldsm4(&bfr[kk][0], b_base + (uint32_t)warp * SS_KB_BYTES + ((((uint32_t)(2 * kk) + b_chunk) ^ (b_row & 7u)) << 4));
      } else {
#pragma unroll
        for (int k2 = 0; k2 < 2; ++k2) {
          uint32_t r4[4];
          ldsm4(r4, b_base + (uint32_t)warp * SS_KB_BYTES + ((((uint32_t)(4 * k2) + b_chunk) ^ (b_row & 7u)) << 4));
          bfr[2 * k2][0] = r4[0]; bfr[2 * k2][1] = r4[1]; bfr[2 * k2 + 1][0] = r4[2]; bfr[2 * k2 + 1][1] = r4[3];
        }
      }
#pragma unroll
      for (int kk = 0; kk < 4; ++kk)
#pragma unroll
        for (int mi = 0; mi < 2; ++mi) {
          const uint32_t r = (uint32_t)(16 * mi + (lane & 7) + 8 * ((lane >> 3) & 1));
          uint32_t a[4];
          ldsm4(a, st + r * 128u + ((((uint32_t)(2 * kk) + a_chunk) ^ (r & 7u)) << 4));
#pragma unroll
          for (int nt = 0; nt < SS_NT; ++nt) mma16816(fa[mi][nt], a, bfr[kk][2 * nt], bfr[kk][2 * nt + 1]);
        }
    }
    float* part = reinterpret_cast<float*>(sm + SO_RING + ((s + 1) % SS_STAGES) * SS_STAGE_BYTES);   // [8 warps][32 o][SS_N]
#pragma unroll
    for (int mi = 0; mi < 2; ++mi)
#pragma unroll
      for (int nt = 0; nt < SS_NT; ++nt)
#pragma unroll
        for (int e = 0; e < 4; ++e)
          part[(warp * SS_LD + 16 * mi + g4 + 8 * (e >> 1)) * SS_N + 8 * nt + 2 * t4 + (e & 1)] = fa[mi][nt][e];
    cw_bar();
    for (int i = tid; i < SS_LD * SS_N; i += SS_CW * 32) {
      const int j = i / SS_LD, o = i % SS_LD, sg = row0 + j;          // consecutive threads: consecutive outputs of a stream
      if (sg < H.n) {
        float a = v_bfin[o];
#pragma unroll
        for (int w = 0; w < SS_CW; ++w) a += part[(w * SS_LD + o) * SS_N + j];
        const float v = fmaf(H.alpha, a, xs[j * SS_LD + o]);
        H.x_out[(size_t)sg * SS_LD + o] = v;
        if (H.prev_latent) {
          H.prev_latent[(size_t)H.slot_ids[sg] * SS_LD + o] = v;
          if (H.commit_out) H.commit_out[(size_t)sg * SS_LD + o] = v;
        }
      }
    }
  }
  SS_TRACE(30);
  tl_mark(tl, 3);
}

}  // namespace

bool flow_head_ss_supported(int flow_dim, int ldim, int depth) {
  return flow_dim == SS_F && ldim == SS_LD && depth >= 1 && depth <= 6;   // 6: the per-feature vectors of all blocks fit in shared memory
}

bool flow_head_ss_prepare(FlowHeadTcParams* p) {
  const FlowHeadParams& h = p->h;
  for (int m = 0; m < 2 * h.depth; ++m) {
    const bf16* w = (m & 1) ? h.blk[m >> 1].w2 : h.blk[m >> 1].w0;
    if (!tc_make_map(&p->maps[m], w, SS_F, SS_F, SS_F, SS_STAGE_ROWS)) return false;
  }
  return tc_make_map(&p->maps[2 * h.depth], h.w_final, h.ldim, SS_F, SS_F, h.ldim);
}

cudaError_t launch_flow_head_ss(const FlowHeadTcParams& p, cudaStream_t st) {
  static DeviceOnce attr_done;
  if (attr_done.first()) {
    cudaError_t e = cudaFuncSetAttribute(flow_head_ss_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)ss_smem_bytes(PTTS_HEAD_MAX_DEPTH > 6 ? 6 : PTTS_HEAD_MAX_DEPTH));
    if (e != cudaSuccess) return e;
  }
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = dim3((p.h.n + SS_N - 1) / SS_N, 1, 1);
  cfg.blockDim = dim3(SS_THREADS, 1, 1);
  cfg.dynamicSmemBytes = ss_smem_bytes(p.h.depth);
  cfg.stream = st;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = pdl_enabled() ? 1 : 0;
  static int trace = -1, calls = 0;
  if (trace < 0) { const char* e = getenv("PTTS_HEAD_TRACE"); trace = (e && e[0] == '1') ? 1 : 0; }
  if (!trace) return cudaLaunchKernelEx(&cfg, flow_head_ss_kernel, p);
  FlowHeadTcParams q = p;
  q.trace = 1;
  cudaError_t e = cudaLaunchKernelEx(&cfg, flow_head_ss_kernel, q);
  cudaStreamCaptureStatus cs = cudaStreamCaptureStatusNone;
  cudaStreamIsCapturing(st, &cs);
  if (e == cudaSuccess && cs == cudaStreamCaptureStatusNone && ++calls == 8) {
    unsigned long long t[64];
    cudaStreamSynchronize(st);
    cudaMemcpyFromSymbol(t, g_ss_trace, sizeof t);
    fprintf(stderr, "[head-ss trace] ns since kernel start:");
    for (int i = 1; i < 31; ++i) if (t[i]) fprintf(stderr, " %d:%lld", i, (long long)(t[i] - t[0]));
    fprintf(stderr, "\n");
  }
  return e;
}

}  // namespace ptts
