// Flow-head kernel, stream-split form (bf16 mode, default): one Euler step of the LSD sampler (mlp.py:86-131,188-215;
// flow_lm.py:32-40) with NO cross-CTA exchange.  Each CTA owns 16 streams and computes the whole 512-wide head for
// them; every CTA streams all the head weights (12 matrices of 512 x 512 bf16 + the final 32 x 512) through a deep TMA
// ring — they are L2-resident after the first CTA touched them.
//
//   orientation   swap-AB: tcgen05.mma with A = weight tile (M = 128 output features, K-major rows of the [N, K]
//                 matrix as stored), B = activation operand (N = 16 streams, K-major), D[feature, stream] in TMEM
//                 (4 feature tiles x 16 columns).  TMEM lane = feature, so one epilogue thread owns a feature of each
//                 tile and 8 of the 16 streams: bias, SiLU, gate, residual and modulation are register-local.
//   LayerNorm     statistics over the 512 features of a stream are a CTA-local reduction (transposing warp butterfly
//                 + one named barrier) instead of the cluster exchange of flow_head_tc.cu.
//   operand       the epilogue writes the next GEMM's B operand (bf16, SWIZZLE_128B k-blocks) straight into shared
//                 memory; two buffers, because tile t's epilogue runs while tiles t+1.. of the same GEMM still read.
//   weight ring   11 stages of 16 KB; the producer runs ahead across GEMM boundaries and starts before the
//                 programmatic-dependent-launch wait (weights depend on nothing).  The fp32 input_proj weight rides
//                 through the first four stages.
//   pipeline      GEMMs are issued tile-major with one commit per feature tile, so the epilogue of tile t overlaps
//                 the MMAs / weight stream of tiles t+1..3.
//
// What bounds it: one SM ingests 512 KB per GEMM from L2 (profiles/probes/l2_ingest_probe.cu measures that rate).
#include <cstdio>
#include <cstdlib>

#include "gemm.cuh"
#include "kernels.cuh"
#include "tc_ptx.cuh"

namespace ptts {

PTTS_TL_ENABLE_FN(head_ss)

namespace {

constexpr int SS_N = 16;                 // streams per CTA = UMMA N
constexpr int SS_F = 512;                // head width
constexpr int SS_TILES = SS_F / 128;     // feature tiles (UMMA M = 128)
constexpr int SS_KB = SS_F / 64;         // k-blocks of 64 (128 B of bf16)
constexpr int SS_EPI_WARPS = 8;          // warps 0..7: TMEM lane quarter = warp % 4, stream half = warp / 4
constexpr int SS_THREADS = (SS_EPI_WARPS + 2) * 32;   // warp 8: TMA producer, warp 9: MMA issuer / TMEM owner
constexpr int SS_STAGES = 11;
constexpr uint32_t SS_STAGE_BYTES = 128 * 128;          // 128 weight rows x 64 bf16
constexpr uint32_t SS_KB_BYTES = SS_N * 128;            // one k-block of the activation operand
constexpr uint32_t SS_OP_BYTES = SS_KB * SS_KB_BYTES;   // 16 KB per operand buffer
constexpr uint32_t SS_TMEM_COLS = 64;
constexpr int SS_LD = 32;                // latent width (checked by the launcher)

constexpr uint32_t SO_OP = 0;                                        // [2][8 kb][16 rows][128 B]
constexpr uint32_t SO_RING = SO_OP + 2 * SS_OP_BYTES;                // [11][128 rows][128 B]
constexpr uint32_t SO_X = SO_RING + SS_STAGES * SS_STAGE_BYTES;      // [16][32] fp32
constexpr uint32_t SO_STAT = SO_X + SS_N * SS_LD * 4;                // [2][8 warps][16] fp32
constexpr uint32_t SO_BAR = SO_STAT + 2 * SS_EPI_WARPS * 16 * 4;
constexpr uint32_t SS_SMEM = SO_BAR + 512 + 1024;

struct SsBars {
  uint64_t wfull[SS_STAGES], wempty[SS_STAGES], accfull[SS_TILES], opfull[2];
  uint32_t tmem_base;
};
static_assert(sizeof(SsBars) <= 512, "barrier block");

__device__ __forceinline__ void wait_bar(uint32_t bar, uint32_t parity) {   // bounded: a protocol error traps
  for (uint32_t spin = 0; spin < (1u << 22); ++spin)
    if (mbar_test_a(bar, parity)) return;
  __trap();
}
__device__ __forceinline__ void epi_bar() { asm volatile("bar.sync 1, 256;" ::: "memory"); }
__device__ __forceinline__ void bulk_g2s(uint32_t dst, const void* src, uint32_t bytes, uint32_t bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
               ::"r"(dst), "l"(src), "r"(bytes), "r"(bar) : "memory");
}
__device__ __forceinline__ void tmem_ld8(uint32_t taddr, float* v) {
  uint32_t r[8];
  asm volatile("tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7])
               : "r"(taddr));
  asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
  for (int i = 0; i < 8; ++i) v[i] = __uint_as_float(r[i]);
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
  const uint32_t accfull_a = smem_u32(&bars.accfull[0]), opfull_a = smem_u32(&bars.opfull[0]);

  pdl_launch_dependents();
  const unsigned tl = tl_begin(6);
  SS_TRACE(0);
  if (tid == 0) {
    for (int s = 0; s < SS_STAGES; ++s) { mbar_init(&bars.wfull[s], 1); mbar_init(&bars.wempty[s], 1); }
    for (int t = 0; t < SS_TILES; ++t) mbar_init(&bars.accfull[t], 1);
    mbar_init(&bars.opfull[0], SS_EPI_WARPS); mbar_init(&bars.opfull[1], SS_EPI_WARPS);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == SS_EPI_WARPS + 1) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&bars.tmem_base)), "n"(SS_TMEM_COLS) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = bars.tmem_base;

  if (warp == SS_EPI_WARPS) {
    // ===== producer: input_proj weight (fp32, 4 stages), then every weight tile of every GEMM in issue order =====
    if (lane == 0) {
      for (int m = 0; m < n_gemm + 1; ++m) asm volatile("prefetch.tensormap [%0];" ::"l"(&P.maps[m]) : "memory");
      uint32_t s = 0, ph = 0;
      for (int i = 0; i < SS_TILES; ++i) {
        mbar_expect_tx_a(wfull_a + 8u * s, SS_STAGE_BYTES);
        bulk_g2s(sm_a + SO_RING + s * SS_STAGE_BYTES, reinterpret_cast<const char*>(H.w_in) + (size_t)i * SS_STAGE_BYTES,
                 SS_STAGE_BYTES, wfull_a + 8u * s);
        ++s;
      }
      for (int m = 0; m <= n_gemm; ++m) {
        const bool fin = m == n_gemm;
        const int tiles = fin ? 1 : SS_TILES;
        const uint32_t bytes = fin ? (uint32_t)SS_LD * 128u : SS_STAGE_BYTES;
        for (int t = 0; t < tiles; ++t)
          for (int kb = 0; kb < SS_KB; ++kb) {
            wait_bar(wempty_a + 8u * s, ph ^ 1u);
            mbar_expect_tx_a(wfull_a + 8u * s, bytes);
            tma_load_2d_a(sm_a + SO_RING + s * SS_STAGE_BYTES, &P.maps[m], wfull_a + 8u * s, kb * 64, t * 128);
            if (++s == SS_STAGES) { s = 0; ph ^= 1u; }
          }
      }
    }
  } else if (warp == SS_EPI_WARPS + 1) {
    // ===== MMA issuer =====
    if (lane == 0) {
      // instruction descriptor: D = f32, A = B = bf16, both K-major, N >> 3 at bit 17, M >> 4 at bit 24
      const uint32_t idesc = (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(SS_N >> 3) << 17) | ((uint32_t)(128 >> 4) << 24);
      uint32_t s = SS_TILES, ph = 0, oph[2] = {0, 0};
      for (int m = 0; m <= n_gemm; ++m) {
        const int buf = m & 1;
        const int tiles = m == n_gemm ? 1 : SS_TILES;
        wait_bar(opfull_a + 8u * buf, oph[buf]);
        oph[buf] ^= 1u;
        tc_fence_after();
        const uint64_t bdesc0 = make_sw128_desc(sm_a + SO_OP + buf * SS_OP_BYTES);
        for (int t = 0; t < tiles; ++t) {
          const uint32_t d_tmem = tmem_base + (uint32_t)(t * SS_N);
          for (int kb = 0; kb < SS_KB; ++kb) {
            wait_bar(wfull_a + 8u * s, ph);
            tc_fence_after();
            const uint64_t adesc = make_sw128_desc(sm_a + SO_RING + s * SS_STAGE_BYTES);
            const uint64_t bdesc = bdesc0 + (uint64_t)((kb * SS_KB_BYTES) >> 4);
#pragma unroll
            for (int k = 0; k < 4; ++k)
              tc_mma_bf16(d_tmem, adesc + (uint64_t)(k * 2), bdesc + (uint64_t)(k * 2), idesc, (kb | k) != 0);
            tc_commit_a(wempty_a + 8u * s);
            if (++s == SS_STAGES) { s = 0; ph ^= 1u; }
          }
          tc_commit_a(accfull_a + 8u * t);
        }
      }
    }
  } else {
    // ===== epilogue warps: thread = feature 32q + lane of every tile, streams 8 hf .. 8 hf + 7 =====
    const int q = warp & 3, hf = warp >> 2;
    const int fl = 32 * q + lane;                       // feature inside a tile
    float* xs = reinterpret_cast<float*>(sm + SO_X);
    float* stat = reinterpret_cast<float*>(sm + SO_STAT);
    const uint32_t tbase = tmem_base + ((uint32_t)(32 * q) << 16) + (uint32_t)(8 * hf);
    // operand element (stream row j of this half, feature f = 128 t + fl): k-block 2t + (q >> 1), 16-byte chunk
    // 4 (q & 1) + lane / 8 XOR-swizzled with the row, element lane % 8
    const uint32_t op_thread = (uint32_t)((q >> 1) * SS_KB_BYTES + hf * 1024 + (lane & 7) * 2);
    const uint32_t op_chunk = (uint32_t)(4 * (q & 1) + (lane >> 3));
    uint32_t accph = 0;
    int stat_buf = 0;

    auto op_store = [&](int buf, int t, int j, float v) {
      const uint32_t off = SO_OP + buf * SS_OP_BYTES + (uint32_t)(2 * t) * SS_KB_BYTES + op_thread + (uint32_t)j * 128u +
                           ((op_chunk ^ (uint32_t)j) << 4);
      *reinterpret_cast<bf16*>(sm + off) = __float2bfloat16_rn(v);
    };
    auto op_publish = [&](int buf) {
      fence_proxy_async();
      __syncwarp();
      if (lane == 0) mbar_arrive_a(opfull_a + 8u * buf);
    };

    pdl_wait();                                         // x_in / mod come from the previous kernels
    tl_mark(tl, 2);
    SS_TRACE(1);
    for (int i = tid; i < SS_N * SS_LD; i += SS_EPI_WARPS * 32) {
      const int r = i / SS_LD;
      xs[i] = (row0 + r < H.n) ? H.x_in[(size_t)row0 * SS_LD + i] : 0.0f;
    }
    bool valid[8];
    const float* mod_row[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      valid[j] = row0 + 8 * hf + j < H.n;
      mod_row[j] = H.mod + (size_t)(valid[j] ? row0 + 8 * hf + j : 0) * H.mod_ld;
    }
    float sh[SS_TILES][8], sc[SS_TILES][8];
    auto load_mod2 = [&](int off) {                     // shift | scale of the block whose modulation starts at `off`
#pragma unroll
      for (int t = 0; t < SS_TILES; ++t)
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          sh[t][j] = valid[j] ? __ldg(mod_row[j] + off + 128 * t + fl) : 0.0f;
          sc[t][j] = valid[j] ? __ldg(mod_row[j] + off + SS_F + 128 * t + fl) : 0.0f;
        }
    };
    load_mod2(0);
    epi_bar();

    // ---- u = input_proj(x): fp32 weight rows wait in ring stages 0..3 (tile t = stage t)
    float u[SS_TILES][8];
#pragma unroll
    for (int t = 0; t < SS_TILES; ++t) {
      const float b = __ldg(H.b_in + 128 * t + fl);
#pragma unroll
      for (int j = 0; j < 8; ++j) u[t][j] = b;
    }
    for (int t = 0; t < SS_TILES; ++t) wait_bar(wfull_a + 8u * t, 0);
#pragma unroll
    for (int c = 0; c < 8; ++c) {
      const int kk = (c + lane) & 7;                    // rotated k order: conflict-free 16-byte reads of 128-byte rows
      float4 xv[8];
#pragma unroll
      for (int j = 0; j < 8; ++j) xv[j] = *reinterpret_cast<const float4*>(xs + (8 * hf + j) * SS_LD + 4 * kk);
#pragma unroll
      for (int t = 0; t < SS_TILES; ++t) {
        const float4 w = *reinterpret_cast<const float4*>(sm + SO_RING + t * SS_STAGE_BYTES + fl * 128 + kk * 16);
#pragma unroll
        for (int j = 0; j < 8; ++j)
          u[t][j] = fmaf(w.x, xv[j].x, fmaf(w.y, xv[j].y, fmaf(w.z, xv[j].z, fmaf(w.w, xv[j].w, u[t][j]))));
      }
    }
    epi_bar();                                          // every thread has read its rows: hand the stages back
    if (tid == 0)
      for (int t = 0; t < SS_TILES; ++t) mbar_arrive_a(wempty_a + 8u * t);
    SS_TRACE(2);

    // LayerNorm statistics of the 16 streams over 512 features: per-thread partials over its 4 features, a
    // transposing butterfly over the 32 lanes (16 values -> 1 per lane), 4 quarter-warps combined through smem.
    auto ln_stats = [&](float* mu, float* rs) {
      float v[16];
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        float s1 = 0.0f, s2 = 0.0f;
#pragma unroll
        for (int t = 0; t < SS_TILES; ++t) { s1 += u[t][j]; s2 = fmaf(u[t][j], u[t][j], s2); }
        v[j] = s1; v[8 + j] = s2;
      }
#pragma unroll
      for (int r = 0; r < 4; ++r) {
        const int half = 8 >> r;
        const bool up = (lane >> r) & 1;
#pragma unroll
        for (int i = 0; i < half; ++i) {
          const float keep = up ? v[i + half] : v[i];
          const float send = up ? v[i] : v[i + half];
          v[i] = keep + __shfl_xor_sync(0xffffffffu, send, 1 << r);
        }
      }
      v[0] += __shfl_xor_sync(0xffffffffu, v[0], 16);
      // lane l (< 16) now holds value index (b0 << 3) | (b1 << 2) | (b2 << 1) | b3 of its bits
      const int idx = ((lane & 1) << 3) | ((lane & 2) << 1) | ((lane & 4) >> 1) | ((lane & 8) >> 3);
      float* sb = stat + stat_buf * (SS_EPI_WARPS * 16);
      if (lane < 16) sb[warp * 16 + idx] = v[0];
      epi_bar();
      float tot = 0.0f;
      if (lane < 16) {
#pragma unroll
        for (int w = 0; w < 4; ++w) tot += sb[(4 * hf + w) * 16 + lane];
      }
      const float e2 = __shfl_down_sync(0xffffffffu, tot, 8) * (1.0f / SS_F);
      const float mean = tot * (1.0f / SS_F);
      const float rstd = rsqrtf(fmaxf(e2 - mean * mean, 0.0f) + 1e-6f);
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        mu[j] = __shfl_sync(0xffffffffu, mean, j);
        rs[j] = __shfl_sync(0xffffffffu, rstd, j);
      }
      stat_buf ^= 1;
    };

    for (int blk = 0; blk < depth; ++blk) {
      const FlowHeadBlock& B = H.blk[blk];
      float mu[8], rs[8];
      ln_stats(mu, rs);
      // g = LN(u) * (1 + scale) + shift -> operand of mlp.0 (buffer 0)
#pragma unroll
      for (int t = 0; t < SS_TILES; ++t) {
        const float lw = __ldg(B.lnw + 128 * t + fl), lb = __ldg(B.lnb + 128 * t + fl);
#pragma unroll
        for (int j = 0; j < 8; ++j)
          op_store(0, t, j, fmaf(fmaf((u[t][j] - mu[j]) * rs[j], lw, lb), 1.0f + sc[t][j], sh[t][j]));
      }
      op_publish(0);
      SS_TRACE(3 + blk * 3);
      // gate of this block and shift | scale of the next one: in flight across both GEMMs
      float gt[SS_TILES][8];
      const int mo = blk * 3 * SS_F;
#pragma unroll
      for (int t = 0; t < SS_TILES; ++t)
#pragma unroll
        for (int j = 0; j < 8; ++j) gt[t][j] = valid[j] ? __ldg(mod_row[j] + mo + 2 * SS_F + 128 * t + fl) : 0.0f;
      load_mod2(mo + 3 * SS_F);
      // mlp.0: h = SiLU(acc + b0) -> operand of mlp.2 (buffer 1), tile by tile
#pragma unroll
      for (int t = 0; t < SS_TILES; ++t) {
        const float b0 = __ldg(B.b0 + 128 * t + fl);
        wait_bar(accfull_a + 8u * t, accph);
        tc_fence_after();
        float a[8];
        tmem_ld8(tbase + (uint32_t)(t * SS_N), a);
#pragma unroll
        for (int j = 0; j < 8; ++j) op_store(1, t, j, silu_call(a[j] + b0));
      }
      accph ^= 1u;
      tc_fence_before();
      op_publish(1);
      SS_TRACE(4 + blk * 3);
      // mlp.2: u += gate * (acc + b2)
#pragma unroll
      for (int t = 0; t < SS_TILES; ++t) {
        const float b2 = __ldg(B.b2 + 128 * t + fl);
        wait_bar(accfull_a + 8u * t, accph);
        tc_fence_after();
        float a[8];
        tmem_ld8(tbase + (uint32_t)(t * SS_N), a);
#pragma unroll
        for (int j = 0; j < 8; ++j) u[t][j] = fmaf(gt[t][j], a[j] + b2, u[t][j]);
      }
      accph ^= 1u;
      tc_fence_before();
      SS_TRACE(5 + blk * 3);
    }
    {  // final layer: LN without affine, modulate, Linear 512 -> ldim, Euler update, commit
      float mu[8], rs[8];
      ln_stats(mu, rs);
#pragma unroll
      for (int t = 0; t < SS_TILES; ++t)
#pragma unroll
        for (int j = 0; j < 8; ++j) op_store(0, t, j, fmaf((u[t][j] - mu[j]) * rs[j], 1.0f + sc[t][j], sh[t][j]));
      op_publish(0);
      if (q == 0) {                                     // TMEM lanes 0..31 = output feature o = lane
        wait_bar(accfull_a, accph);
        tc_fence_after();
        float a[8];
        tmem_ld8(tbase, a);
        const float bf = __ldg(H.b_final + lane);
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          const int sg = row0 + 8 * hf + j;
          if (valid[j]) {
            const float o = fmaf(H.alpha, a[j] + bf, xs[(8 * hf + j) * SS_LD + lane]);
            H.x_out[(size_t)sg * SS_LD + lane] = o;
            if (H.prev_latent) {
              H.prev_latent[(size_t)H.slot_ids[sg] * SS_LD + lane] = o;
              if (H.commit_out) H.commit_out[(size_t)sg * SS_LD + lane] = o;
            }
          }
        }
        tc_fence_before();
      }
    }
  }
  SS_TRACE(30);
  __syncthreads();
  tl_mark(tl, 3);
  if (warp == SS_EPI_WARPS + 1) {
    tc_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "n"(SS_TMEM_COLS) : "memory");
  }
}

}  // namespace

bool flow_head_ss_supported(int flow_dim, int ldim, int depth) {
  return flow_dim == SS_F && ldim == SS_LD && depth >= 1 && depth <= PTTS_HEAD_MAX_DEPTH;
}

bool flow_head_ss_prepare(FlowHeadTcParams* p) {
  const FlowHeadParams& h = p->h;
  for (int m = 0; m < 2 * h.depth; ++m) {
    const bf16* w = (m & 1) ? h.blk[m >> 1].w2 : h.blk[m >> 1].w0;
    if (!tc_make_map(&p->maps[m], w, SS_F, SS_F, SS_F, 128)) return false;
  }
  return tc_make_map(&p->maps[2 * h.depth], h.w_final, h.ldim, SS_F, SS_F, h.ldim);
}

cudaError_t launch_flow_head_ss(const FlowHeadTcParams& p, cudaStream_t st) {
  static DeviceOnce attr_done;
  if (attr_done.first()) {
    cudaError_t e = cudaFuncSetAttribute(flow_head_ss_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)SS_SMEM);
    if (e != cudaSuccess) return e;
  }
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = dim3((p.h.n + SS_N - 1) / SS_N, 1, 1);
  cfg.blockDim = dim3(SS_THREADS, 1, 1);
  cfg.dynamicSmemBytes = SS_SMEM;
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
