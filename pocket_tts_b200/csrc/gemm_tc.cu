// tcgen05 / TMEM / TMA GEMM for sm_100a (bf16 operands, fp32 accumulate in tensor memory).
//
//   D[m, n] = epilogue( sum_k A[m, k] * W[n, k] )
//
// Warp roles (320 threads): warp 0 = TMA producer, warp 1 = TMEM allocator + tcgen05.mma issuer
// (both walk their loops as whole warps, one lane issues), warps 2..9 = epilogue (tcgen05.ld ->
// registers -> fused epilogue -> global; two warps per TMEM lane quarter).  Operand tiles are
// [rows][64 bf16] = 128-byte rows in the canonical K-major SWIZZLE_128B layout, written by TMA and
// read by the tensor core through shared-memory matrix descriptors; a ring of `stages` buffers is
// handed back and forth with mbarriers (full: TMA complete_tx, empty: tcgen05.commit).
//
// Two orientations:
//   SWAP = false : UMMA M (128 TMEM lanes) = activation rows, UMMA N = BN weight rows.
//                  Large-M problems (Mimi transformer, SEANet convolutions as overlapping-row
//                  GEMMs: the A tensor map simply has row stride C_in < K).
//   SWAP = true  : UMMA M = 128 weight rows, UMMA N = all streams (<= 256).  Skinny decode
//                  GEMMs (M = number of streams): the weight matrix is streamed exactly once,
//                  split along K over a thread-block cluster (gridDim.z = cluster size S) so that
//                  many SMs pull HBM.  Every CTA pushes its partial accumulator tile into its
//                  peers' shared memory (st.shared::cluster), CTA r owning features
//                  [r*128/S, (r+1)*128/S): one cluster barrier, then each CTA sums the S slabs in
//                  a fixed order (bit-stable, no atomics, no memset) and runs the full epilogue.
//
// Weight tiles do not depend on the previous kernel of the frame step, so the producer issues
// them before `griddepcontrol.wait`: with programmatic dependent launch the weight stream of
// GEMM i+1 starts while GEMM i is still in its epilogue.
#include <cuda.h>

#include <algorithm>
#include <cstdio>
#include <cstdlib>
#include <map>
#include <mutex>
#include <tuple>

#include "gemm.cuh"
#include "tc_ptx.cuh"

namespace ptts {

PTTS_TL_ENABLE_FN(gemm_tc)

constexpr int BM_TC = 128;   // TMEM lanes / UMMA M
constexpr int BK_TC = 64;    // bf16 elements per 128-byte swizzle row
constexpr int MAX_STAGES = 8;
constexpr int TC_THREADS = 64 + 256;   // producer warp, MMA warp, 8 epilogue warps

struct TcParams {
  int M, N;            // logical output rows (activation rows) and features
  int num_kb;          // K / 64
  int kb_per_split;    // k-blocks handled by one blockIdx.z
  int stages;
  int cluster;         // SWAP split-K: CTAs per cluster along z (partials reduced through DSMEM)
  const bf16* w_base;  // SWAP: weight matrix [N, w_ld] for the L2 prefetch of this CTA's slice (nullptr: none)
  int w_ld;
  KvPrefetchJob pf;    // SWAP: K/V rows of the following decode attention -> L2 (base == nullptr: none)
  int dry_epilogue;    // SWAP: epilogue warps run one storeless pass of the slab epilogue while the mainloop runs (opt-in probe, PTTS_DRY_EPI=1)
  int slab1;           // SWAP, cluster == 1: still go through the slab epilogue (transpose in this CTA's shared memory,
                       // row-contiguous 16-byte stores) instead of the column epilogue (one 4-byte store per stream row)
  long long* trace;    // debug: per-event SM clock stamps of CTA (0,0,0) (PTTS_TC_TRACE=1)
};
#define TC_TRACE(slot) do { if (P.trace && blockIdx.x == 0 && blockIdx.y == 0 && blockIdx.z == 0 && lane == 0) P.trace[slot] = clock64(); } while (0)

// developer trace of CTA 0 (profiles/probes/gemm_rows_bench.cu defines TC_ROWS_TRACE and the buffer): SM clock stamps per tile
#ifdef TC_ROWS_TRACE
__device__ long long g_rows_trace[16 + 8 * 32];
__device__ long long g_epi_trace[16 * 32];
__device__ int g_epi_tile;
#define EPI_TRACE(k) do { if (blockIdx.x == 0 && threadIdx.x == 64 && g_epi_tile < 32) g_epi_trace[g_epi_tile * 16 + (k)] = clock64(); } while (0)
#define ROWS_TRACE(t, k) do { if (blockIdx.x == 0 && lane == 0 && (t) < 32) g_rows_trace[16 + (t) * 8 + (k)] = clock64(); } while (0)
#else
#define ROWS_TRACE(t, k) do { } while (0)
#define EPI_TRACE(k) do { } while (0)
#endif
// Row-major epilogue (SWAP = false).  TMEM hands every thread one output ROW, but global memory
// wants consecutive lanes on consecutive COLUMNS: each epilogue warp therefore stages its
// 32 rows x CH columns chunk in shared memory (padded, conflict free both ways) and re-reads it
// with CH/8 lanes per row, 8 columns per lane, so that bias / gate / residual loads and the
// out / out2 stores are row-contiguous 16-byte accesses.  Per-row addresses (the RowMaps) are
// computed once per row into shared memory.
struct RowMeta {
  long long out, res, out2, pad;   // element offsets, out < 0: row not written (32 bytes: two 16-byte shared accesses)
};
constexpr int EPI_STAGE_FLOATS = 32 * 36;                     // per warp
constexpr size_t EPI_SMEM_BYTES = 8 * EPI_STAGE_FLOATS * 4 + BM_TC * sizeof(RowMeta);

template <int BN, typename TO>
__device__ __forceinline__ void epilogue_rows(const GemmEpilogue& ep, int M, int N, uint32_t taddr, int row0,
                                              int col0, int q, int half, int lane, uint8_t* scratch) {
  constexpr int CH = BN >= 32 ? 32 : 16;
  constexpr int NCH = BN / CH;
  constexpr int LD = CH + 4;
  constexpr int LPR = CH / 8;          // lanes per row when re-reading
  constexpr int RPI = 32 / LPR;        // rows per iteration
  // shared memory by 32-bit address (LDS / STS; through the generic `scratch` pointer these were LD.E / ST.E)
  const uint32_t scratch_a = smem_u32(scratch);
  const uint32_t meta_a = scratch_a + 8u * EPI_STAGE_FLOATS * 4u;
  const uint32_t stage_a = scratch_a + (uint32_t)(q + 4 * half) * (EPI_STAGE_FLOATS * 4u);
  if (half == 0) {
    const int rl = q * 32 + lane, r = row0 + rl;
    RowMeta mt{-1, 0, 0, 0};
    if (r < M) {
      const int ob = r / ep.out_map.rows_per_batch, ot = r - ob * ep.out_map.rows_per_batch;
      if (ot < ep.out_map.valid_rows) {
        mt.out = ob * ep.out_map.batch_stride + (long long)ot * ep.out_map.ld;
        if (ep.res) {
          const int rb = r / ep.res_map.rows_per_batch, rt = r - rb * ep.res_map.rows_per_batch;
          mt.res = rb * ep.res_map.batch_stride + (long long)rt * ep.res_map.ld;
        }
        if (ep.out2) {
          const int b2 = r / ep.out2_map.rows_per_batch, t2 = r - b2 * ep.out2_map.rows_per_batch;
          mt.out2 = b2 * ep.out2_map.batch_stride + (long long)t2 * ep.out2_map.ld;
        }
      }
    }
    sts2l(meta_a + (uint32_t)rl * 32u, mt.out, mt.res);
    sts2l(meta_a + (uint32_t)rl * 32u + 16u, mt.out2, 0);
  }
  EPI_TRACE(0);
  asm volatile("bar.sync 1, 256;" ::: "memory");   // the 8 epilogue warps
  EPI_TRACE(1);
  const int sub = lane / LPR, cg = lane % LPR;
#pragma unroll 1
  for (int ci = half; ci < NCH; ci += 2) {
    const int c = ci * CH, n0 = col0 + c;
    if (n0 >= N) break;   // uniform
    {
      float v[CH];
      tmem_ld<CH>(taddr + c, v);   // warp-collective
      EPI_TRACE(2 + 3 * (ci >> 1));
#pragma unroll
      for (int j = 0; j < CH; j += 4) sts4f(stage_a + (uint32_t)(lane * LD + j) * 4u, make_float4(v[j], v[j + 1], v[j + 2], v[j + 3]));
    }
    __syncwarp();
    EPI_TRACE(3 + 3 * (ci >> 1));
    const int n = n0 + cg * 8;
    float bias[8], cs[8];
    if (ep.bias) row_load<8>(ep.bias + n, bias);
    if (ep.colscale) row_load<8>(ep.colscale + n, cs);
#pragma unroll
    for (int it = 0; it < LPR; ++it) {
      const int rl = q * 32 + it * RPI + sub;
      RowMeta mt;
      lds2l(meta_a + (uint32_t)rl * 32u, mt.out, mt.res);
      if (mt.out >= 0) {
        float x[8], t[8];
        {
          const uint32_t xa = stage_a + (uint32_t)((it * RPI + sub) * LD + cg * 8) * 4u;
          const float4 x0 = lds4f(xa), x1 = lds4f(xa + 16u);
          x[0] = x0.x; x[1] = x0.y; x[2] = x0.z; x[3] = x0.w; x[4] = x1.x; x[5] = x1.y; x[6] = x1.z; x[7] = x1.w;
        }
        if (ep.bias) {
#pragma unroll
          for (int j = 0; j < 8; ++j) x[j] += bias[j];
        }
        act_inplace<8>(x, ep.act_pre);
        if (ep.gate) {
          row_load<8>(ep.gate + (long long)(row0 + rl) * ep.gate_ld + n, t);
#pragma unroll
          for (int j = 0; j < 8; ++j) x[j] *= t[j];
        }
        if (ep.colscale) {
#pragma unroll
          for (int j = 0; j < 8; ++j) x[j] *= cs[j];
        }
        if (ep.alpha != 1.0f) {
#pragma unroll
          for (int j = 0; j < 8; ++j) x[j] *= ep.alpha;
        }
        if (ep.res) {
          row_load<8>(reinterpret_cast<const TO*>(ep.res) + mt.res + n, t);
#pragma unroll
          for (int j = 0; j < 8; ++j) x[j] += t[j];
        }
        if (ep.out2) {
          long long pad;
          lds2l(meta_a + (uint32_t)rl * 32u + 16u, mt.out2, pad);
          row_store<8>(reinterpret_cast<TO*>(ep.out2) + mt.out2 + n, x);
        }
        act_inplace<8>(x, ep.act_post);
        row_store<8>(reinterpret_cast<TO*>(ep.out) + mt.out + n, x);
      }
    }
    __syncwarp();
    EPI_TRACE(4 + 3 * (ci >> 1));
  }
}

// Swapped epilogue without split-K (cluster == 1): this thread owns feature n, TMEM columns are
// streams m.  out / res / out2 / gate are plain row-major [M, ld] (checked on the host).
template <int BN, typename TO>
__device__ __forceinline__ void epilogue_cols(const GemmEpilogue& ep, const TcParams& P, uint32_t taddr, int n,
                                              int m_base, int half, const float2* ln_stat) {
  constexpr int CH = BN >= 32 ? 32 : 16;
  constexpr int NCH = BN / CH;
  const bool n_ok = n < P.N;
  const float bias = (ep.bias && n_ok) ? ep.bias[n] : 0.0f;
  const float scale = ep.alpha * ((ep.colscale && n_ok) ? ep.colscale[n] : 1.0f);
  TO* out = reinterpret_cast<TO*>(ep.out) + n;
  const TO* res = ep.res ? reinterpret_cast<const TO*>(ep.res) + n : nullptr;
  TO* out2 = ep.out2 ? reinterpret_cast<TO*>(ep.out2) + n : nullptr;
  const float* gate = ep.gate ? ep.gate + n : nullptr;
#pragma unroll 1
  for (int ci = half; ci < NCH; ci += 2) {
    const int c = ci * CH, m0 = m_base + c;
    if (m0 >= P.M) break;   // uniform
    float v[CH];
    tmem_ld<CH>(taddr + c, v);
    if (n_ok) {
      const int cnt = min(CH, P.M - m0);
      if (ep.ln_part_in) {
        const float lc1 = ep.ln_c1[n], lc2 = ep.ln_c2[n];
#pragma unroll
        for (int j = 0; j < CH; ++j)
          if (j < cnt) { const float2 st = ln_stat[m0 + j]; v[j] = fmaf(st.y, v[j] - st.x * lc1, lc2); }
      }
#pragma unroll
      for (int j = 0; j < CH; ++j) v[j] += bias;
      act_inplace<CH>(v, ep.act_pre);
#pragma unroll
      for (int j = 0; j < CH; ++j) {
        if (j < cnt) {
          const long long m = m0 + j;
          float x = v[j];
          if (gate) x *= gate[m * ep.gate_ld];
          x *= scale;
          if (res) x += to_f(res[m * ep.res_map.ld]);
          v[j] = x;
          if (out2) out2[m * ep.out2_map.ld] = from_f<TO>(x);
        }
      }
      act_inplace<CH>(v, ep.act_post);
#pragma unroll
      for (int j = 0; j < CH; ++j)
        if (j < cnt) out[(long long)(m0 + j) * ep.out_map.ld] = from_f<TO>(v[j]);
    }
  }
}

// Swapped orientation, split-K over a cluster of S CTAs.  Slab layout in every CTA's shared
// memory: slab[s][m][f'] (s = sending rank, m = stream < BN, f' < F = 128/S owned features).
// Phase 1 (push): the thread holding feature f = owner*F + f' of this CTA's partial tile stores its
// 32 stream values into the owner's slab[rank].  Lanes hold consecutive f', so a warp store covers
// 128/F owners with F*4-byte contiguous segments.
template <int BN>
__device__ __forceinline__ void partial_push(const TcParams& P, uint32_t taddr, const float* slab, int S, int rank,
                                             int f, int half) {
  constexpr int CH = BN >= 32 ? 32 : 16;
  constexpr int NCH = BN / CH;
  const int F = BM_TC / S;
  const int owner = f / F, fo = f - owner * F;
  const uint32_t dst0 = dsmem_addr(smem_u32(slab) + (uint32_t)((rank * BN) * F + fo) * 4u, (uint32_t)owner);
  const uint32_t mstride = (uint32_t)F * 4u;
#pragma unroll 1
  for (int ci = half; ci < NCH; ci += 2) {
    const int c = ci * CH;
    if (c >= P.M) break;   // uniform: no streams in this chunk
    float v[CH];
    tmem_ld<CH>(taddr + c, v);
    const uint32_t dst = dst0 + (uint32_t)c * mstride;
#pragma unroll
    for (int j = 0; j < CH; ++j) dsmem_st(dst + (uint32_t)j * mstride, v[j]);
  }
}

// Phase 2 (after the cluster barrier): sum the S slabs of this CTA in rank order and run the full
// epilogue on features [rank*F, rank*F + F) of the tile with row-contiguous 16-byte accesses.
// per-feature vectors of the slab epilogue (bias, LayerScale * alpha, folded-LayerNorm c1 / c2): weights, so they are
// loaded before the accumulator wait instead of after the cluster barrier (a cold HBM round trip each)
struct SlabVecs {
  float4 bias, cs, lc1, lc2;
};
__device__ __forceinline__ SlabVecs slab_vecs_load(const GemmEpilogue& ep, const TcParams& P, int S, int rank, int row0, int et) {
  const int F = BM_TC / S, FV = F >> 2;
  const int n = row0 + rank * F + (et % FV) * 4;
  SlabVecs v;
  v.bias = make_float4(0.f, 0.f, 0.f, 0.f); v.cs = make_float4(1.f, 1.f, 1.f, 1.f); v.lc1 = v.bias; v.lc2 = v.bias;
  if (n < P.N) {
    if (ep.bias) v.bias = ld4f(ep.bias + n);
    if (ep.colscale) v.cs = ld4f(ep.colscale + n);
    if (ep.ln_part_in) { v.lc1 = ld4f(ep.ln_c1 + n); v.lc2 = ld4f(ep.ln_c2 + n); }
  }
  v.cs.x *= ep.alpha; v.cs.y *= ep.alpha; v.cs.z *= ep.alpha; v.cs.w *= ep.alpha;
  return v;
}

template <int BN, typename TO>
__device__ __forceinline__ void epilogue_slabs(const GemmEpilogue& ep, const TcParams& P, const float* slab, int S,
                                               int rank, int row0, int et, const float2* ln_stat, const SlabVecs& sv, unsigned tl_s = ~0u,
                                               bool dry = false) {
  const int F = BM_TC / S, FV = F >> 2;
  const int fv = et % FV, mstep = 256 / FV;
  const int n = row0 + rank * F + fv * 4;
  if (n >= P.N) return;
  const float4 bias = sv.bias, cs = sv.cs, lc1 = sv.lc1, lc2 = sv.lc2;
  const unsigned grp_mask = 0xFu << ((threadIdx.x & 31) & ~3);   // the 4 lanes that share a row and a 16-feature block
  const float* src = slab + fv * 4;
  for (int m = et / FV; m < P.M; m += mstep) {
    float v[4] = {bias.x, bias.y, bias.z, bias.w};
    for (int s = 0; s < S; ++s) {
      const float4 t = ld4f(src + (s * BN + m) * F);
      v[0] += t.x; v[1] += t.y; v[2] += t.z; v[3] += t.w;
    }
    if (ep.ln_part_in) {   // folded LayerNorm of the operand row (mean, rstd from the producer's partials)
      const float2 st = ln_stat[m];
      v[0] = fmaf(st.y, v[0] - st.x * lc1.x, lc2.x); v[1] = fmaf(st.y, v[1] - st.x * lc1.y, lc2.y);
      v[2] = fmaf(st.y, v[2] - st.x * lc1.z, lc2.z); v[3] = fmaf(st.y, v[3] - st.x * lc1.w, lc2.w);
    }
    act_inplace<4>(v, ep.act_pre);
    if (ep.gate) {
      const float4 g = ld4f(ep.gate + (long long)m * ep.gate_ld + n);
      v[0] *= g.x; v[1] *= g.y; v[2] *= g.z; v[3] *= g.w;
    }
    v[0] *= cs.x; v[1] *= cs.y; v[2] *= cs.z; v[3] *= cs.w;
    if (ep.res) {
      const float4 r4 = load4(reinterpret_cast<const TO*>(ep.res) + (long long)m * ep.res_map.ld + n);
      v[0] += r4.x; v[1] += r4.y; v[2] += r4.z; v[3] += r4.w;
    }
    if (ep.out2 && !dry) store4(reinterpret_cast<TO*>(ep.out2) + (long long)m * ep.out2_map.ld + n, make_float4(v[0], v[1], v[2], v[3]));
    if (ep.ln_part_out) {   // producer side of a folded LayerNorm: bf16 copy + (sum, M2) over this 16-feature block
      if (!dry) store4(ep.ln_raw + (long long)m * P.N + n, make_float4(v[0], v[1], v[2], v[3]));
      float sum = (v[0] + v[1]) + (v[2] + v[3]);
      sum += __shfl_xor_sync(grp_mask, sum, 1);
      sum += __shfl_xor_sync(grp_mask, sum, 2);
      const float mu = sum * (1.0f / 16.0f);
      const float d0 = v[0] - mu, d1 = v[1] - mu, d2 = v[2] - mu, d3 = v[3] - mu;
      float m2 = fmaf(d0, d0, fmaf(d1, d1, fmaf(d2, d2, d3 * d3)));
      m2 += __shfl_xor_sync(grp_mask, m2, 1);
      m2 += __shfl_xor_sync(grp_mask, m2, 2);
      if ((fv & 3) == 0 && !dry) ep.ln_part_out[(long long)(n >> 4) * P.M + m] = make_float2(sum, m2);   // [block][row]: full-sector runs
    }
    act_inplace<4>(v, ep.act_post);
    // dry pass (before the accumulator exists): same instructions and the same residual / gate loads on garbage
    // slab contents, nothing stored - it pulls the epilogue's code into the instruction cache and its operands into
    // L1 while the mainloop runs (the first real pass cost 1.6 us in every GEMM of the chain, timeline)
    if (dry) break;
    store4(reinterpret_cast<TO*>(ep.out) + (long long)m * ep.out_map.ld + n, make_float4(v[0], v[1], v[2], v[3]));
    if (tl_s != ~0u) { tl_mark(tl_s, 5); tl_s = ~0u; }   // timeline: first row pass of the slab epilogue stored
  }
}

// Consumer side of a folded LayerNorm: the 8 epilogue warps combine the per-block (sum, M2) partials of every
// operand row into (mean, rstd) in shared memory while the mainloop runs (parallel-variance formula).
__device__ __forceinline__ void ln_combine(const GemmEpilogue& ep, int M, float2* ln_stat, int et) {
  const int np = ep.ln_nparts;
  const float inv_d = 1.0f / (16.0f * (float)np);
  const unsigned mask = 0xFu << ((threadIdx.x & 31) & ~3);
  const int j = et & 3;
  for (int m = et >> 2; m < M; m += 64) {
    const float2* pr = ep.ln_part_in + m;               // partial i of row m at pr[i * M]
    float tot = 0.0f, q = 0.0f, mu;
    if (np == 64) {   // d_model 1024: 16 partials per lane, all loads in flight at once
      float2 v[16];
#pragma unroll
      for (int i = 0; i < 16; ++i) v[i] = pr[(long long)(16 * j + i) * M];
#pragma unroll
      for (int i = 0; i < 16; ++i) tot += v[i].x;
      tot += __shfl_xor_sync(mask, tot, 1);
      tot += __shfl_xor_sync(mask, tot, 2);
      mu = tot * inv_d;
#pragma unroll
      for (int i = 0; i < 16; ++i) { const float dm = v[i].x * (1.0f / 16.0f) - mu; q += v[i].y + 16.0f * dm * dm; }
    } else {
      for (int i = j; i < np; i += 4) tot += pr[(long long)i * M].x;
      tot += __shfl_xor_sync(mask, tot, 1);
      tot += __shfl_xor_sync(mask, tot, 2);
      mu = tot * inv_d;
      for (int i = j; i < np; i += 4) { const float2 p = pr[(long long)i * M]; const float dm = p.x * (1.0f / 16.0f) - mu; q += p.y + 16.0f * dm * dm; }
    }
    q += __shfl_xor_sync(mask, q, 1);
    q += __shfl_xor_sync(mask, q, 2);
    if (j == 0) ln_stat[m] = make_float2(mu, rsqrtf(q * inv_d + ep.ln_eps));
  }
  asm volatile("bar.sync 2, 256;" ::: "memory");   // the 8 epilogue warps
}

// Register budget: the skinny (SWAP) form belongs to the latency-bound FlowLM chain and must fit on an SM BESIDE one CTA
// of the persistent row-major kernel of the Mimi branch (320 threads x 88 registers): 320 x 112 + 320 x 88 = 64 000 of
// the SM's 65 536.  Left to itself ptxas took 168 (no spills at 112), which made co-residency impossible whatever the
// shared-memory budgets said.  The row-major form needs 72; 96 allows two CTAs per SM.
#ifndef PTTS_SKINNY_MAXREG
#define PTTS_SKINNY_MAXREG 112   // -DPTTS_SKINNY_MAXREG=168 rebuilds the old allocation for A/B runs (PTTS_NATIVE_LIB)
#endif
template <int BN, bool SWAP, typename TO>
__global__ void __maxnreg__(SWAP ? PTTS_SKINNY_MAXREG : 96)
gemm_tc_kernel(const __grid_constant__ CUtensorMap tm_row, const __grid_constant__ CUtensorMap tm_col,
               TcParams P, GemmEpilogue ep) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  // carve: [stages][A 16 KB][B BN*128 B] | slab (cluster split-K only)
  constexpr uint32_t A_BYTES = BM_TC * 128, B_BYTES = BN * 128, STAGE_BYTES = A_BYTES + B_BYTES;
  uint8_t* tiles = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  float* slab = reinterpret_cast<float*>(tiles + (size_t)P.stages * STAGE_BYTES);
  __shared__ __align__(8) uint64_t full_bar[MAX_STAGES];
  __shared__ __align__(8) uint64_t empty_bar[MAX_STAGES];
  __shared__ __align__(8) uint64_t tmem_full_bar;
  __shared__ uint32_t tmem_base_smem;
  __shared__ float2 ln_stat[SWAP ? (BN < 128 ? BN : 128) : 1];   // folded LayerNorm: (mean, rstd) per operand row (M <= 128: a 256-row table would push the 256-stream GEMMs, whose slab is 128 KB, from 2 ring stages to 1)

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (warp == 2) TC_TRACE(4);
  pdl_launch_dependents();               // the next kernel may start its own prologue / weight stream
  __shared__ unsigned tl_slot;
  const unsigned tl = tl_begin((1 << 28) | (P.N << 8) | P.num_kb);
  if (threadIdx.x == 0) tl_slot = tl;
  const int row0 = blockIdx.x * BM_TC;   // UMMA-M operand row offset
  const int col0 = blockIdx.y * BN;      // UMMA-N operand row offset
  const int kb0 = blockIdx.z * P.kb_per_split;
  const int nkb = min(P.kb_per_split, P.num_kb - kb0);
  const int n_pre = min(P.stages, nkb);  // stages filled before the set-up barrier
  constexpr uint32_t TMEM_COLS = BN < 32 ? 32 : BN;
  const bool clustered = SWAP && (P.cluster > 1 || P.slab1);
  // weights are the UMMA-M operand in the swapped orientation, the UMMA-N operand otherwise
  const CUtensorMap* tm_w = SWAP ? &tm_row : &tm_col;
  const CUtensorMap* tm_a = SWAP ? &tm_col : &tm_row;
  const int w_row = SWAP ? row0 : col0, a_row = SWAP ? col0 : row0;
  const uint32_t w_off = SWAP ? 0u : A_BYTES, a_off = SWAP ? A_BYTES : 0u;

  if (SWAP && P.w_base && nkb > n_pre) {
    // the ring holds only n_pre k-blocks before the producer kernel is done: pull the rest of this CTA's weight
    // slice into L2 now (weights depend on nothing), so the later ring passes are L2 hits instead of HBM round trips
    const int lines = nkb - n_pre;                       // 128-byte lines per weight row still to come
    const char* wb = reinterpret_cast<const char*>(P.w_base + (size_t)row0 * P.w_ld + (size_t)(kb0 + n_pre) * BK_TC);
    const int rows = min(BM_TC, P.N - row0);
    for (int idx = threadIdx.x; idx < rows * lines; idx += TC_THREADS) {
      const int r = idx / lines, l = idx - r * lines;
      asm volatile("prefetch.global.L2 [%0];" ::"l"(wb + (size_t)r * P.w_ld * 2 + (size_t)l * 128));
    }
  }
  if (SWAP && P.pf.base && warp >= 2) {
    // side job (KvPrefetchJob): one (stream, k|v, head) region per epilogue warp at a time
    const int G = (int)(gridDim.x * gridDim.z) * 8, wid = ((int)(blockIdx.z * gridDim.x + blockIdx.x)) * 8 + (warp - 2);
    const int R = P.pf.n * P.pf.regions;
    for (int r = wid; r < R; r += G) {
      const int i = r / P.pf.regions, q = r - i * P.pf.regions;
      const int slot = P.pf.slot_ids[i];
      const int bytes = P.pf.offsets[slot] * P.pf.row_bytes;
      const char* p = P.pf.base + (long long)slot * P.pf.slot_stride + (long long)q * P.pf.region_stride;
      for (int off = lane * 128; off < bytes; off += 32 * 128) asm volatile("prefetch.global.L2 [%0];" ::"l"(p + off));
    }
  }
  if (warp == 0 && lane == 0) {
    asm volatile("prefetch.tensormap [%0];" ::"l"(&tm_row) : "memory");
    asm volatile("prefetch.tensormap [%0];" ::"l"(&tm_col) : "memory");
    for (int s = 0; s < P.stages; ++s) { mbar_init(&full_bar[s], 1); mbar_init(&empty_bar[s], 1); }
    mbar_init(&tmem_full_bar, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    // first ring pass: weight tiles first (independent of the previous kernel), then, once the
    // previous kernel's results are visible, the activation tiles of the same stages
    const uint32_t t0a = smem_u32(tiles), f0a = smem_u32(&full_bar[0]);
    for (int i = 0; i < n_pre; ++i) {
      mbar_expect_tx_a(f0a + 8u * i, STAGE_BYTES);
      tma_load_2d_a(t0a + i * STAGE_BYTES + w_off, tm_w, f0a + 8u * i, (kb0 + i) * BK_TC, w_row);
    }
    pdl_wait();
    for (int i = 0; i < n_pre; ++i)
      tma_load_2d_a(t0a + i * STAGE_BYTES + a_off, tm_a, f0a + 8u * i, (kb0 + i) * BK_TC, a_row);
  }
  if (warp == 1) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&tmem_base_smem)), "n"(TMEM_COLS) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  pdl_wait();                            // everything below may read the previous kernel's outputs
  tl_mark(tl, 2);
  if (clustered) cluster_arrive_relaxed();   // "this CTA runs": peers may write its slabs after the matching wait
  const uint32_t tmem_base = tmem_base_smem;
  if (warp == 2) TC_TRACE(0);

  const uint32_t tiles_a = smem_u32(tiles), full_a = smem_u32(&full_bar[0]), empty_a = smem_u32(&empty_bar[0]);
  SlabVecs sv{};
  if (warp == 0) {
    // ===== TMA producer: the whole warp walks the ring, one lane issues.  Stage / phase are
    // tracked incrementally: a runtime i % stages costs more than the two TMA issues. =====
    uint32_t s = 0, ph = 1;   // i = n_pre is the first k-block of ring pass 1 (only reached when n_pre == stages)
    int kcoord = (kb0 + n_pre) * BK_TC;
    for (int i = n_pre; i < nkb; ++i) {
      mbar_wait_a(empty_a + 8u * s, ph ^ 1u);
      TC_TRACE(8 + i * 4 + 0);
      if (lane == 0) {
        const uint32_t st = tiles_a + s * STAGE_BYTES, fb = full_a + 8u * s;
        mbar_expect_tx_a(fb, STAGE_BYTES);
        tma_load_2d_a(st + w_off, tm_w, fb, kcoord, w_row);
        tma_load_2d_a(st + a_off, tm_a, fb, kcoord, a_row);
      }
      __syncwarp();
      kcoord += BK_TC;
      if (++s == (uint32_t)P.stages) { s = 0; ph ^= 1u; }
    }
    if (clustered) cluster_wait_acquire();
  } else if (warp == 1) {
    // ===== MMA issuer: whole warp waits, one lane issues =====
    // instruction descriptor: D=f32, A=B=bf16, both K-major, N>>3 at bit 17, M>>4 at bit 24
    const uint32_t idesc = (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(BN >> 3) << 17) | ((uint32_t)(BM_TC >> 4) << 24);
    const uint64_t adesc0 = make_sw128_desc(tiles_a), bdesc0 = make_sw128_desc(tiles_a + A_BYTES);
    const uint32_t tfull_a = smem_u32(&tmem_full_bar);
    uint32_t s = 0, ph = 0;
    for (int i = 0; i < nkb; ++i) {
      mbar_wait_a(full_a + 8u * s, ph);
      tc_fence_after();
      if (i == 0 && lane == 0 && blockIdx.x + blockIdx.y + blockIdx.z == 0) tl_mark(tl_slot, 4);   // first operands landed
      TC_TRACE(8 + i * 4 + 1);
      if (lane == 0) {
        const uint64_t soff = (uint64_t)((s * STAGE_BYTES) >> 4);   // stage offset in the 16-byte address field
        const uint64_t adesc = adesc0 + soff, bdesc = bdesc0 + soff;
#pragma unroll
        for (int k = 0; k < BK_TC / 16; ++k) {
          // advance 16 bf16 = 32 bytes inside the swizzled 128-byte row: +2 in 16-byte units
          tc_mma_bf16(tmem_base, adesc + (uint64_t)(k * 2), bdesc + (uint64_t)(k * 2), idesc, (i | k) != 0);
        }
        tc_commit_a(empty_a + 8u * s);   // frees the smem stage when these MMAs retire
        if (i == nkb - 1) tc_commit_a(tfull_a);    // accumulator complete
      }
      TC_TRACE(8 + i * 4 + 2);
      __syncwarp();
      if (++s == (uint32_t)P.stages) { s = 0; ph ^= 1u; }
    }
    if (nkb == 0 && lane == 0) tc_commit_a(tfull_a);
    if (clustered) cluster_wait_acquire();
  } else {
    // ===== epilogue warps 2..9: TMEM lane quarter = warp % 4, two warps share a quarter and
    // split its column chunks =====
    const int q = warp & 3, half = (warp - 2) >> 2;
    if (warp == 2) TC_TRACE(1);
    if (SWAP && ep.ln_part_in) ln_combine(ep, P.M, ln_stat, (int)threadIdx.x - 64);
    if (clustered) {
      sv = slab_vecs_load(ep, P, P.cluster, (int)blockIdx.z, row0, (int)threadIdx.x - 64);
      if (P.dry_epilogue)
        epilogue_slabs<BN, TO>(ep, P, slab, P.cluster, (int)blockIdx.z, row0, (int)threadIdx.x - 64, ln_stat, sv, ~0u, true);
    }
    const bool tl_me = warp == 2 && lane == 0 && blockIdx.x + blockIdx.y + blockIdx.z == 0;
    if (tl_me) tl_mark(tl_slot, 5);                  // epilogue warps ready (LayerNorm statistics combined)
    mbar_wait(&tmem_full_bar, 0);
    tc_fence_after();
    if (tl_me) tl_mark(tl_slot, 6);                  // accumulator complete
    if (warp == 2) TC_TRACE(2);
    const int r = row0 + q * 32 + lane;             // UMMA-M index owned by this thread
    const uint32_t taddr = tmem_base + ((uint32_t)(q * 32) << 16);
    if (clustered) {
      cluster_wait_acquire();                        // every peer CTA has started
      partial_push<BN>(P, taddr, slab, P.cluster, (int)blockIdx.z, q * 32 + lane, half);
    } else if (nkb > 0) {
      if (SWAP) {
        epilogue_cols<BN, TO>(ep, P, taddr, r, col0, half, ln_stat);
        if (tl_me) tl_mark(tl_slot, 7);              // column epilogue of warp 2 done
      }
      else epilogue_rows<BN, TO>(ep, P.M, P.N, taddr, row0, col0, q, half, lane, tiles);
    }
    tc_fence_before();
    if (warp == 2) TC_TRACE(3);
  }
  if (clustered) {
    cluster_arrive_release();                 // this CTA's pushes are done ...
    cluster_wait_acquire();                   // ... and so are everybody else's into this CTA
    if (warp == 2 && lane == 0 && blockIdx.x + blockIdx.y + blockIdx.z == 0) tl_mark(tl_slot, 7);   // partial slabs exchanged
    if (warp == 2) TC_TRACE(5);
    if (warp >= 2)
      epilogue_slabs<BN, TO>(ep, P, slab, P.cluster, (int)blockIdx.z, row0, (int)threadIdx.x - 64, ln_stat, sv,
                             (threadIdx.x == 64 && !ep.ln_part_in && blockIdx.x + blockIdx.y + blockIdx.z == 0) ? tl_slot : ~0u);
    if (warp == 2) TC_TRACE(6);
  }
  __syncthreads();
  tl_mark(tl, 3);
  if (warp == 1) {
    tc_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "n"(TMEM_COLS) : "memory");
  }
}

// ---------------------------------------------------------------------------
// Persistent row-major GEMM (Mimi transformer, SEANet convolutions, prefill): one CTA per SM
// walks a static round-robin list of 128 x BN output tiles.  The smem ring keeps running across
// tiles, the accumulator is double-buffered in TMEM (2 x BN columns) so the epilogue of tile i
// overlaps the tcgen05.mma stream of tile i+1, and set-up (barriers, TMEM allocation, tensor-map
// fetch) is paid once per SM instead of once per tile.  One resident CTA of ~110-180 KB leaves the
// rest of the SM (shared memory, TMEM columns) to the kernels of the concurrent FlowLM branch.
struct TcRowsParams {
  int M, N, num_kb, stages;
  int tiles_n, n_tiles;
};


template <int BN, typename TO>
__global__ void __launch_bounds__(TC_THREADS, 1)
gemm_tc_rows_kernel(const __grid_constant__ CUtensorMap tm_a, const __grid_constant__ CUtensorMap tm_w,
                    TcRowsParams P, GemmEpilogue ep) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  constexpr uint32_t A_BYTES = BM_TC * 128, B_BYTES = BN * 128, STAGE_BYTES = A_BYTES + B_BYTES;
  uint8_t* tiles = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  uint8_t* scratch = tiles + (size_t)P.stages * STAGE_BYTES;   // epilogue staging + row metadata
  __shared__ __align__(8) uint64_t full_bar[MAX_STAGES];
  __shared__ __align__(8) uint64_t empty_bar[MAX_STAGES];
  __shared__ __align__(8) uint64_t tfull_bar[2];
  __shared__ __align__(8) uint64_t tempty_bar[2];
  __shared__ uint32_t tmem_base_smem;
  constexpr uint32_t TMEM_COLS = 2 * BN < 32 ? 32 : 2 * BN;

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  pdl_launch_dependents();
  if (warp == 0 && lane == 0) {
    asm volatile("prefetch.tensormap [%0];" ::"l"(&tm_a) : "memory");
    asm volatile("prefetch.tensormap [%0];" ::"l"(&tm_w) : "memory");
    for (int s = 0; s < P.stages; ++s) { mbar_init(&full_bar[s], 1); mbar_init(&empty_bar[s], 1); }
    for (int b = 0; b < 2; ++b) { mbar_init(&tfull_bar[b], 1); mbar_init(&tempty_bar[b], 8); }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 1) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&tmem_base_smem)), "n"(TMEM_COLS) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  pdl_wait();
  const uint32_t tmem_base = tmem_base_smem;
  const uint32_t tiles_a = smem_u32(tiles), full_a = smem_u32(&full_bar[0]), empty_a = smem_u32(&empty_bar[0]);
  const uint32_t tfull_a = smem_u32(&tfull_bar[0]), tempty_a = smem_u32(&tempty_bar[0]);
#ifdef TC_ROWS_TRACE
  if (blockIdx.x == 0 && threadIdx.x == 0) g_rows_trace[0] = clock64();
#endif

  if (warp == 0) {
    // ===== TMA producer =====
    uint32_t s = 0, ph = 0, tt = 0;
    for (int tile = blockIdx.x; tile < P.n_tiles; tile += gridDim.x, ++tt) {
      const int mt = tile / P.tiles_n, nt = tile - mt * P.tiles_n;
      const int row0 = mt * BM_TC, col0 = nt * BN;
      for (int kb = 0; kb < P.num_kb; ++kb) {
        mbar_wait_a(empty_a + 8u * s, ph ^ 1u);
        if (kb == 0) ROWS_TRACE(tt, 0); else if (kb == P.num_kb - 1) ROWS_TRACE(tt, 1);
        if (lane == 0) {
          const uint32_t st = tiles_a + s * STAGE_BYTES, fb = full_a + 8u * s;
          mbar_expect_tx_a(fb, STAGE_BYTES);
          tma_load_2d_a(st + A_BYTES, &tm_w, fb, kb * BK_TC, col0);
          tma_load_2d_a(st, &tm_a, fb, kb * BK_TC, row0);
        }
        __syncwarp();
        if (++s == (uint32_t)P.stages) { s = 0; ph ^= 1u; }
      }
    }
  } else if (warp == 1) {
    // ===== MMA issuer =====
    const uint32_t idesc = (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(BN >> 3) << 17) | ((uint32_t)(BM_TC >> 4) << 24);
    const uint64_t adesc0 = make_sw128_desc(tiles_a), bdesc0 = make_sw128_desc(tiles_a + A_BYTES);
    uint32_t s = 0, ph = 0, t = 0;
    for (int tile = blockIdx.x; tile < P.n_tiles; tile += gridDim.x, ++t) {
      const uint32_t buf = t & 1u, use = t >> 1;
      mbar_wait_a(tempty_a + 8u * buf, (use & 1u) ^ 1u);   // epilogue drained this accumulator (first use: passes)
      tc_fence_after();
      ROWS_TRACE(t, 2);
      const uint32_t d_tmem = tmem_base + buf * BN;
      for (int kb = 0; kb < P.num_kb; ++kb) {
        mbar_wait_a(full_a + 8u * s, ph);
        tc_fence_after();
        if (kb == 0) ROWS_TRACE(t, 3); else if (kb == P.num_kb - 1) ROWS_TRACE(t, 4);
        if (lane == 0) {
          const uint64_t soff = (uint64_t)((s * STAGE_BYTES) >> 4);
          const uint64_t adesc = adesc0 + soff, bdesc = bdesc0 + soff;
#pragma unroll
          for (int k = 0; k < BK_TC / 16; ++k)
            tc_mma_bf16(d_tmem, adesc + (uint64_t)(k * 2), bdesc + (uint64_t)(k * 2), idesc, (kb | k) != 0);
          tc_commit_a(empty_a + 8u * s);
          if (kb == P.num_kb - 1) tc_commit_a(tfull_a + 8u * buf);
        }
        __syncwarp();
        if (++s == (uint32_t)P.stages) { s = 0; ph ^= 1u; }
      }
    }
  } else {
    // ===== epilogue warps 2..9 =====
    const int q = warp & 3, half = (warp - 2) >> 2;
    uint32_t t = 0;
    for (int tile = blockIdx.x; tile < P.n_tiles; tile += gridDim.x, ++t) {
      const uint32_t buf = t & 1u, use = t >> 1;
      const int mt = tile / P.tiles_n, nt = tile - mt * P.tiles_n;
      mbar_wait_a(tfull_a + 8u * buf, use & 1u);
      tc_fence_after();
      if (warp == 2) ROWS_TRACE(t, 5);
      const uint32_t taddr = tmem_base + buf * BN + ((uint32_t)(q * 32) << 16);
      asm volatile("bar.sync 1, 256;" ::: "memory");   // previous tile's row metadata no longer in use
      epilogue_rows<BN, TO>(ep, P.M, P.N, taddr, mt * BM_TC, nt * BN, q, half, lane, scratch);
      tc_fence_before();
      __syncwarp();
      if (warp == 2) ROWS_TRACE(t, 6);
#ifdef TC_ROWS_TRACE
      if (blockIdx.x == 0 && threadIdx.x == 64) ++g_epi_tile;
#endif
      if (lane == 0) mbar_arrive_a(tempty_a + 8u * buf);
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "n"(TMEM_COLS) : "memory");
  }
}

// --------------------------------------------------------------------------- host side
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static EncodeTiledFn get_encode() {
  static EncodeTiledFn fn = nullptr;
  static std::once_flag once;
  std::call_once(once, [] {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult qres;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &qres) == cudaSuccess &&
        qres == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<EncodeTiledFn>(p);
  });
  return fn;
}

// 2-D bf16 tensor map: inner dim K (contiguous), outer dim `rows` with stride `ld` elements
// (ld < K is allowed: overlapping rows = streaming convolution windows), box = 64 x box_rows.
bool tc_make_map(CUtensorMap* out, const void* base, long long rows, int K, long long ld, int box_rows) {
  using Key = std::tuple<const void*, long long, int, long long, int>;
  static std::map<Key, CUtensorMap> cache;
  static std::mutex mu;
  std::lock_guard<std::mutex> lock(mu);
  Key key{base, rows, K, ld, box_rows};
  auto it = cache.find(key);
  if (it != cache.end()) { *out = it->second; return true; }
  EncodeTiledFn enc = get_encode();
  if (!enc) return false;
  cuuint64_t dims[2] = {(cuuint64_t)K, (cuuint64_t)rows};
  cuuint64_t strides[1] = {(cuuint64_t)ld * 2};
  cuuint32_t box[2] = {(cuuint32_t)BK_TC, (cuuint32_t)box_rows};
  cuuint32_t estr[2] = {1, 1};
  CUtensorMap m;
  CUresult r = enc(&m, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(base), dims, strides, box, estr,
                   CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                   CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) return false;
  cache[key] = m;
  *out = m;
  return true;
}

constexpr size_t TC_SMEM_LIMIT = 224 * 1024;
constexpr size_t TC_CORESIDENT_SMEM = 108 * 1024;   // two CTAs per SM
constexpr size_t TC_ROWS_SMEM = 112 * 1024;         // persistent row-major kernel: leaves ~85 KB per SM to the other branch

template <int BN, bool SWAP, typename TO>
static cudaError_t launch_cfg(const CUtensorMap& tr, const CUtensorMap& tc, const TcParams& P, const GemmEpilogue& ep,
                              dim3 grid, bool pdl, cudaStream_t st) {
  const bool clustered = SWAP && (P.cluster > 1 || P.slab1);
  size_t smem = (size_t)P.stages * (BM_TC * 128 + BN * 128) + (clustered ? (size_t)BN * BM_TC * sizeof(float) : 0);
  if (!SWAP) smem = std::max(smem, EPI_SMEM_BYTES);   // the epilogue staging reuses the (drained) pipeline buffers
  smem += 1024;
  // PTTS_MIMI_HALF=1 (probe): row-major GEMMs (the Mimi branch) take more than half an SM's shared memory, i.e. one CTA
  // per SM, so that a CTA of the latency-critical FlowLM branch always finds room beside it
  static const bool half_sm = [] { const char* e = getenv("PTTS_MIMI_HALF"); return e && e[0] == '1'; }();
  if (!SWAP && half_sm) smem = std::max<size_t>(smem, 118 * 1024);
  static DeviceOnce attr_done;
  auto kern = gemm_tc_kernel<BN, SWAP, TO>;
  if (attr_done.first()) {
    // opt-in limit is 227 KB for static + dynamic together: the static part (barriers, LayerNorm statistics) varies with BN
    cudaFuncAttributes fa{};
    cudaError_t e = cudaFuncGetAttributes(&fa, kern);
    if (e != cudaSuccess) return e;
    const int dyn_max = std::min((int)TC_SMEM_LIMIT + 1024, 232448 - (int)fa.sharedSizeBytes);
    e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, dyn_max);
    if (e != cudaSuccess) return e;
    prefer_max_smem_carveout(reinterpret_cast<const void*>(kern));
  }
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = grid;
  cfg.blockDim = dim3(TC_THREADS, 1, 1);
  cfg.dynamicSmemBytes = smem;
  cfg.stream = st;
  cudaLaunchAttribute attr[2];
  int na = 0;
  if (clustered && P.cluster > 1) {
    attr[na].id = cudaLaunchAttributeClusterDimension;
    attr[na].val.clusterDim.x = 1;
    attr[na].val.clusterDim.y = 1;
    attr[na].val.clusterDim.z = P.cluster;
    ++na;
  }
  if (pdl) {
    attr[na].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[na].val.programmaticStreamSerializationAllowed = 1;
    ++na;
  }
  cfg.attrs = attr;
  cfg.numAttrs = na;
  return cudaLaunchKernelEx(&cfg, kern, tr, tc, P, ep);
}

// debug trace (PTTS_TC_TRACE=1): 8 + 4*kb slots of clock64 stamps, dumped by tc_trace_dump()
static long long* g_trace = nullptr;
static long long* tc_trace_buffer() {
  static int enabled = -1;
  if (enabled < 0) { const char* e = getenv("PTTS_TC_TRACE"); enabled = (e && e[0] == '1') ? 1 : 0; }
  if (!enabled) return nullptr;
  if (!g_trace) { cudaMalloc(&g_trace, 1024 * sizeof(long long)); }
  cudaMemset(g_trace, 0, 1024 * sizeof(long long));
  return g_trace;
}
void tc_trace_dump(const char* what) {
  if (!g_trace) return;
  long long h[1024];
  cudaDeviceSynchronize();
  cudaMemcpy(h, g_trace, sizeof h, cudaMemcpyDeviceToHost);
  const long long t0 = h[0];
  fprintf(stderr, "[tc trace %s] kernel_start %lld | setup_done 0 | tmem_full %lld epi/push_done %lld | cluster_barrier %lld slab_epilogue_done %lld\n",
          what, h[4] - t0, h[2] - t0, h[3] - t0, h[5] ? h[5] - t0 : 0, h[6] ? h[6] - t0 : 0);
  for (int i = 0; i < 64 && h[8 + i * 4 + 1]; ++i)
    fprintf(stderr, "  kb %2d: producer_got_empty %7lld  mma_got_full %7lld  mma_issued %7lld\n", i, h[8 + i * 4] - t0,
            h[8 + i * 4 + 1] - t0, h[8 + i * 4 + 2] - t0);
}

template <int BN, typename TO>
static cudaError_t launch_rows(const CUtensorMap& ta, const CUtensorMap& tw, const TcRowsParams& P, const GemmEpilogue& ep,
                               int grid, bool pdl, cudaStream_t st) {
  const size_t smem = (size_t)P.stages * (BM_TC * 128 + BN * 128) + EPI_SMEM_BYTES + 1024;
  static DeviceOnce attr_done;
  auto kern = gemm_tc_rows_kernel<BN, TO>;
  if (attr_done.first()) {
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)TC_SMEM_LIMIT + 1024);
    if (e != cudaSuccess) return e;
  }
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = dim3(grid, 1, 1);
  cfg.blockDim = dim3(TC_THREADS, 1, 1);
  cfg.dynamicSmemBytes = smem;
  cfg.stream = st;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = pdl ? 1 : 0;
  return cudaLaunchKernelEx(&cfg, kern, ta, tw, P, ep);
}

static int pick_bn(int n) {
  if (n <= 16) return 16;
  if (n <= 32) return 32;
  if (n <= 64) return 64;
  if (n <= 128) return 128;
  return 256;
}

// a_rows_alloc: number of addressable A rows (for the tensor-map extent; rows beyond are zero-filled)
cudaError_t gemm_tc_launch(const GemmProblem& p, int out_type, long long a_rows_alloc, int n_sms, bool pdl,
                           cudaStream_t st) {
  if (p.M <= 0 || p.N <= 0) return cudaSuccess;
  if (p.K % BK_TC != 0) return cudaErrorInvalidValue;
  const bool plain_a = p.a_map.rows_per_batch >= (1 << 30);
  if (!plain_a && p.a_map.batch_stride != (long long)p.a_map.rows_per_batch * p.a_map.ld) return cudaErrorInvalidValue;
  const bool plain_out = p.ep.out_map.rows_per_batch >= (1 << 30) && p.ep.res_map.rows_per_batch >= (1 << 30) &&
                         p.ep.out2_map.rows_per_batch >= (1 << 30);
  const bool swap = plain_a && plain_out && p.M <= 256;
  if (!swap && (p.ep.ln_part_in || p.ep.ln_part_out)) return cudaErrorInvalidValue;   // folded LayerNorm: skinny path only
  if (!swap && p.N % 32 != 0 && p.N != 16) return cudaErrorInvalidValue;   // row-major epilogue works in 32-column chunks
  TcParams P{};
  P.M = p.M; P.N = p.N; P.num_kb = p.K / BK_TC;
  P.cluster = 1;
  P.trace = tc_trace_buffer();
  CUtensorMap ta, tw;
  GemmEpilogue ep = p.ep;
  if (swap) {
    const int bn = pick_bn(p.M);
    if (!tc_make_map(&tw, p.W, p.N, p.K, p.K, BM_TC)) return cudaErrorUnknown;
    if (!tc_make_map(&ta, p.A, a_rows_alloc, p.K, p.a_map.ld, bn)) return cudaErrorUnknown;
    static const bool w_prefetch = [] { const char* e = getenv("PTTS_W_PREFETCH"); return !(e && e[0] == '0'); }();
    P.w_base = w_prefetch ? reinterpret_cast<const bf16*>(p.W) : nullptr;
    P.w_ld = p.K;
    P.pf = p.kv_prefetch;
    static const bool dry_epi = [] { const char* e = getenv("PTTS_DRY_EPI"); return e && e[0] == '1'; }();   // measured: no gain (the 1.6 us first pass is not instruction fetch)
    P.dry_epilogue = dry_epi ? 1 : 0;
    const int tiles = (p.N + BM_TC - 1) / BM_TC;
    // split K over a cluster of S CTAs (S in 1,2,4,8) so that many SMs stream the weight matrix
    int splits = 1;
    while (splits < 8 && splits * 2 <= p.max_splits && splits * 2 <= P.num_kb && tiles * splits * 2 <= n_sms) splits *= 2;
    // wide matrices (80 tiles of the AdaLN modulation) run the column epilogue for ~10 us (timeline); a 2-way split with the
    // slab epilogue halves CTA 0's time but its 160 CTAs no longer fit one wave beside the early-launched kernels: 13 -> 22 us
    // on the chain (measured), so it stays opt-in
    static const bool wide_split = [] { const char* e = getenv("PTTS_WIDE_SPLIT"); return e && e[0] == '1'; }();
    if (wide_split && splits == 1 && p.max_splits >= 2 && tiles <= n_sms && P.num_kb >= 4) splits = 2;
    if (p.N % 4 != 0) splits = 1;   // the slab epilogue works on 4-feature vectors
    P.kb_per_split = (P.num_kb + splits - 1) / splits;
    if ((P.num_kb + P.kb_per_split - 1) / P.kb_per_split != splits) { splits = 1; P.kb_per_split = P.num_kb; }
    P.cluster = splits;
    if (ep.ln_part_out && (splits == 1 || p.N % 16 != 0 || !ep.ln_raw)) return cudaErrorInvalidValue;   // partials come from the slab epilogue
    if (ep.ln_part_in && (ep.ln_nparts * 16 != p.K || !ep.ln_c1 || !ep.ln_c2 || p.M > 128)) return cudaErrorInvalidValue;
    static const bool cols_slab = [] { const char* e = getenv("PTTS_COLS_SLAB"); return !(e && e[0] == '0'); }();
    P.slab1 = (cols_slab && splits == 1 && p.N % 4 == 0 && !ep.ln_part_out) ? 1 : 0;
    const size_t stage = (size_t)(BM_TC + bn) * 128, slab = (splits > 1 || P.slab1) ? (size_t)bn * BM_TC * sizeof(float) : 0;
    // keep a CTA under ~half an SM's shared memory when possible: kernels of the two graph branches
    // (FlowLM chain || Mimi chain) and consecutive waves can then share an SM
    const size_t budget = (stage + slab <= TC_CORESIDENT_SMEM) ? TC_CORESIDENT_SMEM : TC_SMEM_LIMIT;
    P.stages = (int)std::max<size_t>(1, std::min<size_t>(std::min(P.kb_per_split, 6), (budget - slab) / stage));
    dim3 grid(tiles, 1, splits);
#define TC_CASE(BNV)                                                                          \
  case BNV:                                                                                   \
    return out_type == 0 ? launch_cfg<BNV, true, float>(tw, ta, P, ep, grid, pdl, st)         \
                         : launch_cfg<BNV, true, bf16>(tw, ta, P, ep, grid, pdl, st);
    switch (bn) {
      TC_CASE(16) TC_CASE(32) TC_CASE(64) TC_CASE(128)
      default: return out_type == 0 ? launch_cfg<256, true, float>(tw, ta, P, ep, grid, pdl, st)
                                    : launch_cfg<256, true, bf16>(tw, ta, P, ep, grid, pdl, st);
    }
#undef TC_CASE
  }
  // row-major: persistent kernel, one CTA per SM over a static tile list.  Tile width from a small
  // cost model (clock trace): ~0.24 us per k-block of tcgen05.mma issue (independent of N) against
  // ~0.5 us + 0.3 us per 64 columns of epilogue, which overlap across tiles; ~3 us to start up.
  const int tiles_m = (p.M + BM_TC - 1) / BM_TC;
  static const bool old_rows = [] { const char* e = getenv("PTTS_TC_ROWS"); return e && e[0] == 'o'; }();
  // measured (ncu, B=64): the persistent kernel wins when a tile has >= 8 k-blocks of mainloop to hide its
  // epilogue behind; thin layers (K <= 384) do better as one tile per CTA with two CTAs per SM
  static const int rows_min_kb = [] { const char* e = getenv("PTTS_TC_ROWS_MINKB"); return e ? atoi(e) : 8; }();
  if (!old_rows && p.N % 32 == 0 && P.num_kb >= rows_min_kb) {
    // at most 128 columns per accumulator (2 x 128 TMEM columns per CTA): a resident 512-column CTA would make
    // every tcgen05.alloc of the concurrent FlowLM branch on that SM wait until it retires
    int bn = std::min(pick_bn(p.N), 128);
    if (bn < 32) bn = 32;
    double best = 1e30;
    for (int cand = bn; cand >= 32; cand >>= 1) {
      const int tiles = tiles_m * ((p.N + cand - 1) / cand);
      const int per_cta = (tiles + n_sms - 1) / n_sms;
      const double epi = 0.5 + 0.3 * cand / 64.0, main = 0.24 * P.num_kb;
      const double t = 3.0 + per_cta * std::max(main, epi) + std::min(main, epi);
      if (t < best - 1e-9) { best = t; bn = cand; }
    }
    TcRowsParams R{};
    R.M = p.M; R.N = p.N; R.num_kb = P.num_kb;
    R.tiles_n = (p.N + bn - 1) / bn;
    R.n_tiles = tiles_m * R.tiles_n;
    const size_t stage = (size_t)(BM_TC + bn) * 128;
    static const size_t rows_smem = [] { const char* e = getenv("PTTS_TC_ROWS_SMEM_KB"); return e ? (size_t)atoi(e) * 1024 : TC_ROWS_SMEM; }();
    const size_t budget = (2 * stage + EPI_SMEM_BYTES <= rows_smem) ? rows_smem : TC_SMEM_LIMIT;
    R.stages = (int)std::max<size_t>(2, std::min<size_t>(6, (budget - EPI_SMEM_BYTES) / stage));
    if (!tc_make_map(&ta, p.A, a_rows_alloc, p.K, p.a_map.ld, BM_TC)) return cudaErrorUnknown;
    if (!tc_make_map(&tw, p.W, p.N, p.K, p.K, bn)) return cudaErrorUnknown;
    const int grid = std::min(R.n_tiles, n_sms);
#define TC_CASE(BNV)                                                                  \
  case BNV:                                                                           \
    return out_type == 0 ? launch_rows<BNV, float>(ta, tw, R, ep, grid, pdl, st)      \
                         : launch_rows<BNV, bf16>(ta, tw, R, ep, grid, pdl, st);
    switch (bn) {
      TC_CASE(32) TC_CASE(64) TC_CASE(128)
      default: return out_type == 0 ? launch_rows<256, float>(ta, tw, R, ep, grid, pdl, st)
                                    : launch_rows<256, bf16>(ta, tw, R, ep, grid, pdl, st);
    }
#undef TC_CASE
  }
  // row-major tile width from a small cost model (measured with the clock trace): a CTA costs
  // ~1.5 us set-up + 0.24 us per k-block (four tcgen05.mma issues + commit, independent of N) +
  // ~2 us epilogue latency + 0.3 us per 64 columns; the grid runs in waves of n_sms CTAs.
  int bn = std::min(pick_bn(p.N), 128);
  {
    double best = 1e30;
    for (int cand = bn; cand >= 32; cand >>= 1) {
      const int ctas = tiles_m * ((p.N + cand - 1) / cand);
      // multi-wave grids run two CTAs per SM (stage ring capped at ~108 KB below): one CTA's
      // epilogue overlaps the other's mainloop, so a "wave" holds 2 * n_sms CTAs at ~1.3x the time
      const bool two = ctas > n_sms;
      const int waves = two ? (ctas + 2 * n_sms - 1) / (2 * n_sms) : 1;
      const double t = waves * (two ? 1.3 : 1.0) * (3.5 + 0.24 * P.num_kb + 0.3 * cand / 64.0);
      if (t < best - 1e-9) { best = t; bn = cand; }
    }
  }
  if (!tc_make_map(&ta, p.A, a_rows_alloc, p.K, p.a_map.ld, BM_TC)) return cudaErrorUnknown;
  if (!tc_make_map(&tw, p.W, p.N, p.K, p.K, bn)) return cudaErrorUnknown;
  P.kb_per_split = P.num_kb;
  {
    const int ctas = tiles_m * ((p.N + bn - 1) / bn);
    (void)ctas;
    const size_t budget = ((size_t)(BM_TC + bn) * 128 <= TC_CORESIDENT_SMEM) ? TC_CORESIDENT_SMEM : TC_SMEM_LIMIT;
    P.stages = (int)std::max<size_t>(1, std::min<size_t>(std::min(P.num_kb, 6), budget / ((size_t)(BM_TC + bn) * 128)));
  }
  dim3 grid(tiles_m, (p.N + bn - 1) / bn, 1);
#define TC_CASE(BNV)                                                                          \
  case BNV:                                                                                   \
    return out_type == 0 ? launch_cfg<BNV, false, float>(ta, tw, P, ep, grid, pdl, st)        \
                         : launch_cfg<BNV, false, bf16>(ta, tw, P, ep, grid, pdl, st);
  switch (bn) {
    TC_CASE(16) TC_CASE(32) TC_CASE(64) TC_CASE(128)
    default: return out_type == 0 ? launch_cfg<256, false, float>(ta, tw, P, ep, grid, pdl, st)
                                  : launch_cfg<256, false, bf16>(ta, tw, P, ep, grid, pdl, st);
  }
#undef TC_CASE
}

}  // namespace ptts
