// Persistent FlowLM frame-step kernel (see flow_mega.h for the phase model).
//
// CTA layout (320 threads): warp 0 = TMA producer, warp 1 = tcgen05.mma issuer, warps 2..9 =
// compute warps (GEMM epilogues out of TMEM, row-wise phases, attention).  One CTA per SM; item i
// of a phase belongs to CTA i % G.  All cross-CTA data goes through global memory (L2): writers
// fence + arrive on the phase's counter, readers acquire it; data written inside the launch is read
// with ld.global.cg / TMA only (never through a possibly stale L1 line).
//
// GEMM orientation: UMMA M = 128 TMEM lanes = stream rows (only `rows_pad` of them are loaded, the
// rest of the M operand aliases the weight tile: those accumulator lanes are never read), UMMA N =
// n_tile weight rows, K in 64-element (128-byte, SWIZZLE_128B) blocks.  A ring stage holds one
// activation block [rows_pad][64] and one weight block [<= 64][64].
#include "flow_mega.h"

#include <cstdio>

#include "tc_ptx.cuh"

namespace ptts {
namespace {

constexpr int MEGA_NCW = 4;                      // compute warps (192 threads / ~100 KB per CTA: an SM can host a Mimi CTA beside it)
constexpr int MEGA_CT = 32 * MEGA_NCW;
constexpr int MEGA_THREADS = 64 + MEGA_CT;
constexpr int MEGA_W_BYTES = 64 * 128;          // weight block of a ring stage
constexpr int MEGA_MAX_STAGES = 12;
constexpr int MEGA_TMEM_COLS = 128;             // two accumulator buffers of 64 columns
constexpr int MEGA_SCRATCH = 1024;               // row-op reduction scratch
constexpr long long MEGA_SPIN_LIMIT = 400000000LL;   // ~0.2 s of SM clocks: a wedged launch aborts instead of hanging

__device__ __forceinline__ unsigned ld_acquire_u32(const unsigned* p) {
  unsigned v;
  asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void red_release_u32(unsigned* p) {
  asm volatile("red.release.gpu.global.add.u32 [%0], 1;" ::"l"(p) : "memory");
}
__device__ __forceinline__ unsigned long long globaltimer_ns() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  return t;
}
__device__ __forceinline__ float4 ldcg4(const float* p) { return __ldcg(reinterpret_cast<const float4*>(p)); }

// bounded spinning: a wait that exceeds MEGA_SPIN_LIMIT clocks records where it was stuck in host-mapped
// memory and traps (the launch fails with an error instead of hanging the device)
struct Watch {
  long long t0;
  unsigned n;
};
__device__ __forceinline__ void watch_poll(Watch& w, const MegaParams& P, unsigned code) {
  if ((++w.n & 63u) != 0) return;
  if (clock64() - w.t0 > MEGA_SPIN_LIMIT) {
    if (P.diag) {
      volatile unsigned* d = P.diag;
      d[1] = blockIdx.x; d[2] = threadIdx.x; d[0] = code;
      __threadfence_system();
    }
    __trap();
  }
}
__device__ __forceinline__ void mbar_wait_bounded(uint32_t bar, uint32_t parity, const MegaParams& P, unsigned code) {
  if (mbar_test_a(bar, parity)) return;
  Watch w{clock64(), 0};
  while (!mbar_test_a(bar, parity)) watch_poll(w, P, code);
}
// all phases [0, need) complete?  (every CTA reaches the phase counters in order).  One lane polls.
__device__ __forceinline__ void grid_wait(const MegaParams& P, int need, int& done_upto, int lane, unsigned code) {
  if (need <= done_upto) return;
  if (lane == 0) {
    const unsigned* c = P.sync + (need - 1);
    Watch w{clock64(), 0};
    while (ld_acquire_u32(c) < (unsigned)P.G) watch_poll(w, P, code);
  }
  __syncwarp();
  done_upto = need;
}

struct Item {
  const MegaGemm* g;
  int tile, split, nkb, w_row, kb0;
};
__device__ __forceinline__ void decode_item(const MegaPhase* ph, int item, Item& it) {
  const MegaGemm* g = &ph->g[0];
  int local = item;
  const int n0 = g->tiles_n * g->splits;
  if (local >= n0) { g = &ph->g[1]; local -= n0; }
  it.g = g;
  it.split = local / g->tiles_n;
  it.tile = local - it.split * g->tiles_n;
  it.nkb = g->kb_per_split;
  it.w_row = it.tile * g->n_tile;
  it.kb0 = it.split * g->kb_per_split;
}

// ------------------------------------------------------------------ TMA producer (warp 0)
struct KCursor {          // walks this CTA's k-blocks through all GEMM phases of the program
  int p, item, kb;
  int nkb, wmap, amap, n_tile, w_row, kb0, rot, kcol;
  bool valid;
};
__device__ __forceinline__ void cursor_load(KCursor& c, const MegaParams& P, bool prefetch_maps) {
  while (c.p < P.n_phases) {
    const MegaPhase* ph = P.phases + c.p;
    if (ph->type == MP_GEMM && c.item < ph->n_items) {   // GEMM phases carry absolute item counts
      Item it;
      decode_item(ph, c.item, it);
      c.nkb = it.nkb; c.wmap = it.g->wmap; c.amap = it.g->amap; c.n_tile = it.g->n_tile;
      // every CTA reads the same activation k-blocks: start each CTA at a different one so that the requests of a
      // phase spread over the L2 slices instead of queueing on one block at a time (any order sums the same set)
      c.w_row = it.w_row; c.kb0 = it.kb0; c.rot = (int)(blockIdx.x % (unsigned)it.nkb); c.kb = 0;
      c.kcol = (c.kb0 + c.rot) * 64;
      c.valid = true;
      if (prefetch_maps && (threadIdx.x & 31) == 0) {
        asm volatile("prefetch.tensormap [%0];" ::"l"(P.maps + c.wmap) : "memory");
        asm volatile("prefetch.tensormap [%0];" ::"l"(P.maps + c.amap) : "memory");
      }
      return;
    }
    ++c.p;
    c.item = blockIdx.x;
  }
  c.valid = false;
}
__device__ __forceinline__ void cursor_next(KCursor& c, const MegaParams& P, bool prefetch_maps) {
  if (++c.kb < c.nkb) {
    int k = c.kb + c.rot;
    if (k >= c.nkb) k -= c.nkb;
    c.kcol = (c.kb0 + k) * 64;
    return;
  }
  c.item += P.G;
  cursor_load(c, P, prefetch_maps);
}

__device__ void producer_role(const MegaParams& P, uint32_t tiles_a, uint32_t full_a, uint32_t empty_a, int lane) {
  const uint32_t stage_bytes = (uint32_t)P.a_bytes + MEGA_W_BYTES;
  KCursor cw, ca;
  cw.p = 0; cw.item = blockIdx.x; cursor_load(cw, P, true);
  ca = cw;
  uint32_t sw = 0, phw = 0, sa = 0;
  long long qw = 0, qa = 0;   // k-blocks whose weight / activation load has been issued
  int done_upto = 0;
  Watch wt{clock64(), 0};
  while (ca.valid) {
    bool progressed = false;
    // weight block of k-block qw: needs only a free ring stage
    const bool wfree = cw.valid && __shfl_sync(0xffffffffu, (int)mbar_probe_a(empty_a + 8u * sw, phw ^ 1u), 0);
    if (wfree) {
      if (lane == 0) {
        const uint32_t fb = full_a + 8u * sw;
        mbar_expect_tx_a(fb, (uint32_t)P.a_bytes + (uint32_t)cw.n_tile * 128u);
        tma_load_2d_a(tiles_a + sw * stage_bytes + (uint32_t)P.a_bytes, P.maps + cw.wmap, fb, cw.kcol, cw.w_row);
      }
      ++qw;
      if (++sw == (uint32_t)P.stages) { sw = 0; phw ^= 1u; }
      cursor_next(cw, P, true);
      progressed = true;
    }
    // activation block of k-block qa: its producers are the phases before ca.p
    if (qa < qw) {
      bool ok = ca.p <= done_upto;
      if (!ok) {
        unsigned cnt = 0;
        if (lane == 0) cnt = ld_acquire_u32(P.sync + (ca.p - 1));
        cnt = __shfl_sync(0xffffffffu, cnt, 0);
        if (cnt >= (unsigned)P.G) {
          done_upto = ca.p;
          ok = true;
          fence_proxy_async();
        }
      }
      if (ok) {
        if (lane == 0)
          tma_load_2d_a(tiles_a + sa * stage_bytes, P.maps + ca.amap, full_a + 8u * sa, ca.kcol, 0);
        ++qa;
        if (++sa == (uint32_t)P.stages) sa = 0;
        cursor_next(ca, P, false);
        progressed = true;
      }
    }
    if (progressed) { wt.n = 0; wt.t0 = clock64(); }
    else watch_poll(wt, P, 0x10000u + (unsigned)ca.p);
    __syncwarp();
  }
}

// ------------------------------------------------------------------ MMA issuer (warp 1)
__device__ void mma_role(const MegaParams& P, uint32_t tiles_a, uint32_t full_a, uint32_t empty_a, uint32_t tfull_a,
                         uint32_t tempty_a, uint32_t tmem_base, int lane) {
  const uint32_t stage_bytes = (uint32_t)P.a_bytes + MEGA_W_BYTES;
  const uint64_t adesc0 = make_sw128_desc(tiles_a), bdesc0 = make_sw128_desc(tiles_a + (uint32_t)P.a_bytes);
  uint32_t s = 0, ph = 0, acc = 0;
  for (int p = 0; p < P.n_phases; ++p) {
    const MegaPhase* phs = P.phases + p;
    if (phs->type != MP_GEMM) continue;
    const int n_items = phs->n_items;
    for (int item = blockIdx.x; item < n_items; item += P.G, ++acc) {
      Item it;
      decode_item(phs, item, it);
      const uint32_t buf = acc & 1u, use = acc >> 1;
      mbar_wait_bounded(tempty_a + 8u * buf, (use & 1u) ^ 1u, P, 0x20000u + (unsigned)p);
      tc_fence_after();
      const uint32_t d_tmem = tmem_base + buf * 64u;
      // instruction descriptor: D=f32, A=B=bf16, both K-major, N>>3 at bit 17, M>>4 at bit 24
      const uint32_t idesc = (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(it.g->n_tile >> 3) << 17) | ((128u >> 4) << 24);
      for (int kb = 0; kb < it.nkb; ++kb) {
        mbar_wait_bounded(full_a + 8u * s, ph, P, 0x30000u + (unsigned)p);
        tc_fence_after();
        if (lane == 0) {
          const uint64_t soff = (uint64_t)((s * stage_bytes) >> 4);
          const uint64_t adesc = adesc0 + soff, bdesc = bdesc0 + soff;
#pragma unroll
          for (int k = 0; k < 4; ++k)
            tc_mma_bf16(d_tmem, adesc + (uint64_t)(k * 2), bdesc + (uint64_t)(k * 2), idesc, (kb | k) != 0);
          tc_commit_a(empty_a + 8u * s);
          if (kb == it.nkb - 1) tc_commit_a(tfull_a + 8u * buf);
        }
        __syncwarp();
        if (++s == (uint32_t)P.stages) { s = 0; ph ^= 1u; }
      }
    }
  }
}

// ------------------------------------------------------------------ compute warps: GEMM epilogue
__device__ __forceinline__ void store16_f32(float* o, const float* v) {
#pragma unroll
  for (int j = 0; j < 16; j += 4) *reinterpret_cast<float4*>(o + j) = make_float4(v[j], v[j + 1], v[j + 2], v[j + 3]);
}
__device__ __forceinline__ void gemm_epilogue(const MegaParams& P, const Item& it, uint32_t taddr, int q, int half, int lane) {
  const MegaGemm* g = it.g;
  const int r = q * 32 + lane;
  const bool row_ok = r < P.n;
  const int n_tile = g->n_tile, mode = g->mode, act = g->act;
  const float* bias = g->bias;
  for (int c = half * 16; c < n_tile; c += 16 * (MEGA_NCW / 4)) {
    float v[16];
    tmem_ld<16>(taddr + (uint32_t)c, v);   // warp-collective
    if (!row_ok) continue;
    const int n0 = it.w_row + c;
    if (mode == MG_PARTIAL) {
      store16_f32(reinterpret_cast<float*>(g->out) + ((size_t)it.split * P.rows_pad + r) * g->nf + n0, v);
    } else {
      if (bias) {
#pragma unroll
        for (int j = 0; j < 16; j += 4) {
          const float4 b4 = *reinterpret_cast<const float4*>(bias + n0 + j);
          v[j] += b4.x; v[j + 1] += b4.y; v[j + 2] += b4.z; v[j + 3] += b4.w;
        }
      }
      act_inplace<16>(v, act);
      if (mode == MG_FULL_F32) store16_f32(reinterpret_cast<float*>(g->out) + (size_t)r * g->out_ld + n0, v);
      else {
        bf16* o = reinterpret_cast<bf16*>(g->out) + (size_t)r * g->out_ld + n0;
        row_store<8>(o, v);
        row_store<8>(o + 8, v + 8);
      }
    }
  }
}

// ------------------------------------------------------------------ compute warps: row-wise phases
// One stream row per CTA: the NCW compute warps split the row's columns, so every load of a row op is issued
// at once (h + up to 8 partial slabs per thread), and reduce through shared memory.
__device__ __forceinline__ void cta_sync_compute() { asm volatile("bar.sync 4, %0;" ::"n"(MEGA_CT) : "memory"); }
__device__ __forceinline__ void cta_sync_phase() { asm volatile("bar.sync 1, %0;" ::"n"(MEGA_CT) : "memory"); }
__device__ __forceinline__ void cta_sync_start() { asm volatile("bar.sync 5, %0;" ::"n"(MEGA_CT) : "memory"); }
__device__ __forceinline__ float block_sum_c(float v, float* red, int cw, int lane) {
  v = warp_sum(v);
  if (lane == 0) red[cw] = v;
  cta_sync_compute();
  float t = 0.0f;
#pragma unroll
  for (int j = 0; j < MEGA_NCW; ++j) t += red[j];
  return t;
}

// NE consecutive output features c0.. of y = W x, W [*, 32] fp32 row-major, x[j] held by lane j of every warp
template <int NE>
__device__ __forceinline__ void dot_rows32(const float* __restrict__ w, int c0, float xl, float* out) {
  float xs[32];
#pragma unroll
  for (int j = 0; j < 32; ++j) xs[j] = __shfl_sync(0xffffffffu, xl, j);
#pragma unroll 2
  for (int e = 0; e < NE; ++e) {
    float4 wr[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) wr[j] = *reinterpret_cast<const float4*>(w + (size_t)(c0 + e) * 32 + j * 4);
    float acc = 0.0f;
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      acc = fmaf(wr[j].x, xs[4 * j], acc);
      acc = fmaf(wr[j].y, xs[4 * j + 1], acc);
      acc = fmaf(wr[j].z, xs[4 * j + 2], acc);
      acc = fmaf(wr[j].w, xs[4 * j + 3], acc);
    }
    out[e] = acc;
  }
}

// h = W_in where(isnan(prev), bos, prev)   (flow_lm.py:121-122).  One warp per (row, 128-feature chunk).
__device__ void row_embed(const MegaParams& P, int item, int lane) {
  const int chunks = P.D / 128;
  const int row = item / chunks, chunk = item - row * chunks;
  const int slot = P.slot_ids[row];
  float xl = 0.0f;
  if (lane < P.Ld) {
    xl = __ldcg(P.prev_latent + (size_t)slot * P.Ld + lane);
    if (isnan(xl)) xl = P.bos[lane];
  }
  const int c0 = chunk * 128 + lane * 4;
  float o[4];
  dot_rows32<4>(P.w_in, c0, xl, o);
  *reinterpret_cast<float4*>(P.h + (size_t)row * P.D + c0) = make_float4(o[0], o[1], o[2], o[3]);
}

// h += sum of partial slabs; out = LayerNorm(h) (bf16)        mimi_transformer.py:45-54
// final variant: out_norm + EOS logit, then the stream's offset advances   flow_lm.py:128-129,153-154
__device__ void row_res_ln(const MegaParams& P, const MegaPhase& ph, int row, int tid_c, int cw, int lane, float* red,
                           bool final_norm) {
  constexpr int J = 1024 / (4 * MEGA_CT);   // float4s per thread (d == 1024)
  const int d = ph.d;
  const size_t ro = (size_t)row * d, slab = (size_t)P.rows_pad * d;
  float4 x[J], t[8][J], wv[J], bv[J];
#pragma unroll
  for (int j = 0; j < J; ++j) x[j] = ldcg4(P.h + ro + (tid_c + MEGA_CT * j) * 4);
#pragma unroll
  for (int s = 0; s < 8; ++s)
    if (s < ph.S) {
#pragma unroll
      for (int j = 0; j < J; ++j) t[s][j] = ldcg4(ph.part + (size_t)s * slab + ro + (tid_c + MEGA_CT * j) * 4);
    }
#pragma unroll
  for (int j = 0; j < J; ++j) {
    wv[j] = *reinterpret_cast<const float4*>(ph.w + (tid_c + MEGA_CT * j) * 4);
    bv[j] = *reinterpret_cast<const float4*>(ph.b + (tid_c + MEGA_CT * j) * 4);
  }
#pragma unroll
  for (int s = 0; s < 8; ++s)
    if (s < ph.S) {
#pragma unroll
      for (int j = 0; j < J; ++j) { x[j].x += t[s][j].x; x[j].y += t[s][j].y; x[j].z += t[s][j].z; x[j].w += t[s][j].w; }
    }
  float sum = 0.0f;
#pragma unroll
  for (int j = 0; j < J; ++j) {
    if (!final_norm && ph.S > 0) *reinterpret_cast<float4*>(P.h + ro + (tid_c + MEGA_CT * j) * 4) = x[j];
    sum += x[j].x + x[j].y + x[j].z + x[j].w;
  }
  const float mu = block_sum_c(sum, red, cw, lane) / (float)d;
  float sq = 0.0f;
#pragma unroll
  for (int j = 0; j < J; ++j) {
    x[j].x -= mu; x[j].y -= mu; x[j].z -= mu; x[j].w -= mu;
    sq += x[j].x * x[j].x + x[j].y * x[j].y + x[j].z * x[j].z + x[j].w * x[j].w;
  }
  const float rstd = rsqrtf(block_sum_c(sq, red + 8, cw, lane) / (float)d + ph.eps);
  float dot = 0.0f;
#pragma unroll
  for (int j = 0; j < J; ++j) {
    const int c = (tid_c + MEGA_CT * j) * 4;
    float4 o;
    o.x = x[j].x * rstd * wv[j].x + bv[j].x;
    o.y = x[j].y * rstd * wv[j].y + bv[j].y;
    o.z = x[j].z * rstd * wv[j].z + bv[j].z;
    o.w = x[j].w * rstd * wv[j].w + bv[j].w;
    store4(reinterpret_cast<bf16*>(ph.out_act) + ro + c, o);
    if (final_norm) {
      const float4 ev = *reinterpret_cast<const float4*>(P.eos_w + c);
      dot += o.x * ev.x + o.y * ev.y + o.z * ev.z + o.w * ev.w;
    }
  }
  if (final_norm) {
    dot = block_sum_c(dot, red + 16, cw, lane);
    if (tid_c == 0) {
      P.eos_logit[row] = dot + P.eos_b[0];
      P.offsets[P.slot_ids[row]] += 1;   // every attention phase of this step has already read it
    }
  }
}

// items [0, n): cemb = cond_embed(c) (+bias), sy = SiLU(cemb + (te(s)+te(t))/2) (mlp.py:203-208), x = noise on the
// first LSD step (flow_lm.py:32-40); items [n, 2n): u = input_proj(x) (mlp.py:202).  Width 512: one float4 per thread.
static_assert(MEGA_CT * 4 == 512, "flow-head row ops assume one float4 per compute thread");
__device__ void row_sy(const MegaParams& P, const MegaPhase& ph, int item, int tid_c, int lane) {
  const int F = P.F;
  const int c = tid_c * 4;
  if (item < P.n) {
    const int row = item;
    const size_t ro = (size_t)row * F + c, slab = (size_t)P.rows_pad * F;
    float4 x;
    if (ph.flag) {
      float4 t[8];
#pragma unroll
      for (int s = 0; s < 8; ++s)
        if (s < ph.S) t[s] = ldcg4(ph.part + (size_t)s * slab + ro);
      x = *reinterpret_cast<const float4*>(P.cond_b + c);
#pragma unroll
      for (int s = 0; s < 8; ++s)
        if (s < ph.S) { x.x += t[s].x; x.y += t[s].y; x.z += t[s].z; x.w += t[s].w; }
      *reinterpret_cast<float4*>(P.cemb + ro) = x;
      if (tid_c < P.Ld / 4) *reinterpret_cast<float4*>(P.x_lat + (size_t)row * P.Ld + tid_c * 4) = ldcg4(P.noise + (size_t)row * P.Ld + tid_c * 4);
    } else {
      x = ldcg4(P.cemb + ro);
    }
    const float4 tc = *reinterpret_cast<const float4*>(P.tcomb + (size_t)ph.idx * F + c);
    store4(reinterpret_cast<bf16*>(ph.out_act) + ro,
           make_float4(silu_f(x.x + tc.x), silu_f(x.y + tc.y), silu_f(x.z + tc.z), silu_f(x.w + tc.w)));
  } else {
    const int row = item - P.n;
    const float* xs = ph.flag ? P.noise : P.x_lat;
    const float xl = lane < P.Ld ? __ldcg(xs + (size_t)row * P.Ld + lane) : 0.0f;
    float o[4];
    dot_rows32<4>(P.w_ip, c, xl, o);
    const float4 b4 = *reinterpret_cast<const float4*>(P.b_ip + c);
    *reinterpret_cast<float4*>(P.u + (size_t)row * F + c) = make_float4(o[0] + b4.x, o[1] + b4.y, o[2] + b4.z, o[3] + b4.w);
  }
}

// pending ResBlock update u += gate * (mlp2 partials + b2) (mlp.py:107-111), then the flow head's LayerNorm
// (mlp.py:49-55) and modulate(x, shift, scale) (mlp.py:16-17) -> bf16 operand of the next GEMM
__device__ void row_lnmod(const MegaParams& P, const MegaPhase& ph, int row, int tid_c, int cw, int lane, float* red) {
  const int F = P.F;
  const int c = tid_c * 4;
  const size_t ro = (size_t)row * F + c, slab = (size_t)P.rows_pad * F, mo = (size_t)row * ph.mod_ld + c;
  float4 x = ldcg4(P.u + ro);
  const float4 sh = ldcg4(ph.shift + mo), sc = ldcg4(ph.scale + mo);
  if (ph.flag) {
    float4 t[8];
#pragma unroll
    for (int s = 0; s < 8; ++s)
      if (s < ph.S) t[s] = ldcg4(ph.part + (size_t)s * slab + ro);
    const float4 gt = ldcg4(ph.gate + mo);
    float4 y = *reinterpret_cast<const float4*>(ph.bias + c);
#pragma unroll
    for (int s = 0; s < 8; ++s)
      if (s < ph.S) { y.x += t[s].x; y.y += t[s].y; y.z += t[s].z; y.w += t[s].w; }
    x.x = fmaf(gt.x, y.x, x.x); x.y = fmaf(gt.y, y.y, x.y);
    x.z = fmaf(gt.z, y.z, x.z); x.w = fmaf(gt.w, y.w, x.w);
    *reinterpret_cast<float4*>(P.u + ro) = x;
  }
  const float mu = block_sum_c(x.x + x.y + x.z + x.w, red, cw, lane) / (float)F;
  const float a0 = x.x - mu, a1 = x.y - mu, a2 = x.z - mu, a3 = x.w - mu;
  const float var = block_sum_c(a0 * a0 + a1 * a1 + a2 * a2 + a3 * a3, red + 8, cw, lane) / (float)F;
  const float den = sqrtf(var + ph.eps);
  float4 o = make_float4(a0 / den, a1 / den, a2 / den, a3 / den);
  if (ph.w) {
    const float4 wv = *reinterpret_cast<const float4*>(ph.w + c), bv = *reinterpret_cast<const float4*>(ph.b + c);
    o.x = o.x * wv.x + bv.x; o.y = o.y * wv.y + bv.y; o.z = o.z * wv.z + bv.z; o.w = o.w * wv.w + bv.w;
  }
  o.x = o.x * (1.0f + sc.x) + sh.x;
  o.y = o.y * (1.0f + sc.y) + sh.y;
  o.z = o.z * (1.0f + sc.z) + sh.z;
  o.w = o.w * (1.0f + sc.w) + sh.w;
  store4(reinterpret_cast<bf16*>(ph.out_act) + ro, o);
}

// Euler step x += (final_layer partials + bias) / lsd (flow_lm.py:36-39); last step: the latent is fed back
__device__ void row_commit(const MegaParams& P, const MegaPhase& ph, int row, int tid_c) {
  const int Ld = P.Ld;
  if (tid_c >= Ld / 4) return;
  const size_t ro = (size_t)row * Ld + tid_c * 4;
  float4 t[8];
#pragma unroll
  for (int s = 0; s < 8; ++s)
    if (s < ph.S) t[s] = ldcg4(ph.part + ((size_t)s * P.rows_pad + row) * Ld + tid_c * 4);
  float4 x = ldcg4(P.x_lat + ro);
  float4 v = *reinterpret_cast<const float4*>(ph.bias + tid_c * 4);
#pragma unroll
  for (int s = 0; s < 8; ++s)
    if (s < ph.S) { v.x += t[s].x; v.y += t[s].y; v.z += t[s].z; v.w += t[s].w; }
  const float a = 1.0f / (float)P.lsd;
  x.x = fmaf(a, v.x, x.x); x.y = fmaf(a, v.y, x.y); x.z = fmaf(a, v.z, x.z); x.w = fmaf(a, v.w, x.w);
  *reinterpret_cast<float4*>(P.x_lat + ro) = x;
  if (ph.flag) {
    *reinterpret_cast<float4*>(P.prev_latent + (size_t)P.slot_ids[row] * Ld + tid_c * 4) = x;
    if (P.commit_out) *reinterpret_cast<float4*>(P.commit_out + ro) = x;
  }
}

// ------------------------------------------------------------------ compute warps: decode attention
// One (stream, head) per WARP (transformer.py:135-158, rope.py:28-58): q/k/v = sum of the QKV partial slabs,
// RoPE at the stream's offset, append k/v, then stream the cached K/V once.  A lane owns a 16-byte slice of a
// K / V row; a warp load covers 4 positions; two batches of 4 K + 4 V loads per lane are kept in flight
// (software double buffering); every 8-lane group keeps an online-softmax state that is merged by shuffles.
__device__ __forceinline__ void bf16x8_to_f32(const uint4& r, float* f) {
  const uint32_t w[4] = {r.x, r.y, r.z, r.w};
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const float2 t = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&w[i]));
    f[2 * i] = t.x; f[2 * i + 1] = t.y;
  }
}
struct SmState {
  float m, l, acc[8];
};
__device__ __forceinline__ void sm_update4(SmState& st, const float* q, const uint4* kr, const uint4* vr, int p_first, int L) {
  float sc[4];
#pragma unroll
  for (int u = 0; u < 4; ++u) {
    float kf[8];
    bf16x8_to_f32(kr[u], kf);
    float a = 0.0f;
#pragma unroll
    for (int e = 0; e < 8; ++e) a = fmaf(q[e], kf[e], a);
    sc[u] = a;
  }
#pragma unroll
  for (int o = 1; o < 8; o <<= 1) {
#pragma unroll
    for (int u = 0; u < 4; ++u) sc[u] += __shfl_xor_sync(0xffffffffu, sc[u], o);
  }
  float mx = st.m;
#pragma unroll
  for (int u = 0; u < 4; ++u) {
    if (p_first + u * 4 >= L) sc[u] = -INFINITY;
    mx = fmaxf(mx, sc[u]);
  }
  if (mx > -INFINITY) {
    const float corr = __expf(st.m - mx);
    st.l *= corr;
#pragma unroll
    for (int e = 0; e < 8; ++e) st.acc[e] *= corr;
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      const float pr = __expf(sc[u] - mx);
      st.l += pr;
      float vf[8];
      bf16x8_to_f32(vr[u], vf);
#pragma unroll
      for (int e = 0; e < 8; ++e) st.acc[e] = fmaf(pr, vf[e], st.acc[e]);
    }
    st.m = mx;
  }
}
__device__ __forceinline__ void kv_load4(uint4* kr, uint4* vr, const bf16* kbase, const bf16* vbase, int p_first, int L) {
#pragma unroll
  for (int u = 0; u < 4; ++u) {
    const int p = p_first + u * 4;
    if (p < L) {
      kr[u] = __ldcg(reinterpret_cast<const uint4*>(kbase + (size_t)p * 64));
      vr[u] = __ldcg(reinterpret_cast<const uint4*>(vbase + (size_t)p * 64));
    } else {
      kr[u] = make_uint4(0, 0, 0, 0);
      vr[u] = make_uint4(0, 0, 0, 0);
    }
  }
}
__device__ __forceinline__ float bf16_round(float x) { return __bfloat162float(__float2bfloat16_rn(x)); }

__device__ void attn_item(const MegaParams& P, const MegaPhase& ph, int item, int slot, int pos, int lane) {
  const int H = P.H, D = P.D;
  const int i = item / H, h = item - i * H;
  const int sub = lane >> 3, ch = lane & 7;
  const size_t slab = (size_t)P.rows_pad * 3 * D;
  const float* row = ph.part + (size_t)i * 3 * D + h * 64 + ch * 8;
  const KvGeom& g = P.kv;
  bf16* kv0 = reinterpret_cast<bf16*>(g.base) + ((((size_t)ph.idx * g.n_slots + slot) * 2) * H + h) * (size_t)g.cap * 64;
  bf16* kbase = kv0 + ch * 8;
  bf16* vbase = kv0 + (size_t)H * g.cap * 64 + ch * 8;
  const int L = pos;   // cached positions [0, pos); the new position is folded in from registers

  // first K/V batch does not depend on this step's projections: issue it before anything else
  uint4 k0[4], v0[4], k1[4], v1[4];
  kv_load4(k0, v0, kbase, vbase, sub, L);

  float q[8], kn[8], vn[8];
  {
    float4 tq[2][2], tk[2][2], tv[2][2];
#pragma unroll
    for (int s = 0; s < 2; ++s)
      if (s < ph.S) {
        const float* r = row + (size_t)s * slab;
        tq[s][0] = ldcg4(r); tq[s][1] = ldcg4(r + 4);
        if (sub == 0) {
          tk[s][0] = ldcg4(r + D); tk[s][1] = ldcg4(r + D + 4);
          tv[s][0] = ldcg4(r + 2 * D); tv[s][1] = ldcg4(r + 2 * D + 4);
        }
      }
#pragma unroll
    for (int e = 0; e < 8; ++e) { q[e] = 0.f; kn[e] = 0.f; vn[e] = 0.f; }
#pragma unroll
    for (int s = 0; s < 2; ++s)
      if (s < ph.S) {
#pragma unroll
        for (int hh = 0; hh < 2; ++hh) {
          q[4 * hh] += tq[s][hh].x; q[4 * hh + 1] += tq[s][hh].y; q[4 * hh + 2] += tq[s][hh].z; q[4 * hh + 3] += tq[s][hh].w;
          if (sub == 0) {
            kn[4 * hh] += tk[s][hh].x; kn[4 * hh + 1] += tk[s][hh].y; kn[4 * hh + 2] += tk[s][hh].z; kn[4 * hh + 3] += tk[s][hh].w;
            vn[4 * hh] += tv[s][hh].x; vn[4 * hh + 1] += tv[s][hh].y; vn[4 * hh + 2] += tv[s][hh].z; vn[4 * hh + 3] += tv[s][hh].w;
          }
        }
      }
#pragma unroll
    for (int e = 0; e < 8; e += 2) {
      float sn, cs;
      sincosf(P.rope_freqs[(ch * 8 + e) >> 1] * (float)pos, &sn, &cs);   // rope.py:42-50
      const float q0 = q[e], q1 = q[e + 1];
      q[e] = (q0 * cs - q1 * sn) * 0.125f;          // 1/sqrt(64) folded into q
      q[e + 1] = (q0 * sn + q1 * cs) * 0.125f;
      const float a0 = kn[e], a1 = kn[e + 1];
      kn[e] = bf16_round(a0 * cs - a1 * sn);        // the cache holds bf16: use the same values now as later steps will read
      kn[e + 1] = bf16_round(a0 * sn + a1 * cs);
      vn[e] = bf16_round(vn[e]);
      vn[e + 1] = bf16_round(vn[e + 1]);
    }
    if (sub == 0) {
      row_store<8>(kbase + (size_t)pos * 64, kn);
      row_store<8>(vbase + (size_t)pos * 64, vn);
    }
  }

  SmState st;
  st.m = -INFINITY; st.l = 0.0f;
#pragma unroll
  for (int e = 0; e < 8; ++e) st.acc[e] = 0.0f;
  const int nb = (L + 15) >> 4;
  for (int b = 0; b < nb; b += 2) {
    if (b + 1 < nb) kv_load4(k1, v1, kbase, vbase, (b + 1) * 16 + sub, L);
    sm_update4(st, q, k0, v0, b * 16 + sub, L);
    if (b + 1 < nb) {
      if (b + 2 < nb) kv_load4(k0, v0, kbase, vbase, (b + 2) * 16 + sub, L);
      sm_update4(st, q, k1, v1, (b + 1) * 16 + sub, L);
    }
  }
  // the new position (registers of the sub == 0 lanes)
  {
    float a = 0.0f;
#pragma unroll
    for (int e = 0; e < 8; ++e) a = fmaf(q[e], kn[e], a);
#pragma unroll
    for (int o = 1; o < 8; o <<= 1) a += __shfl_xor_sync(0xffffffffu, a, o);
    if (sub == 0) {
      const float mx = fmaxf(st.m, a);
      const float corr = __expf(st.m - mx), pr = __expf(a - mx);
      st.l = st.l * corr + pr;
#pragma unroll
      for (int e = 0; e < 8; ++e) st.acc[e] = fmaf(pr, vn[e], st.acc[e] * corr);
      st.m = mx;
    }
  }
  // merge the four lane groups
#pragma unroll
  for (int off = 8; off < 32; off <<= 1) {
    const float m2 = __shfl_xor_sync(0xffffffffu, st.m, off), l2 = __shfl_xor_sync(0xffffffffu, st.l, off);
    const float mm = fmaxf(st.m, m2);
    const float c1 = st.m == -INFINITY ? 0.0f : __expf(st.m - mm), c2 = m2 == -INFINITY ? 0.0f : __expf(m2 - mm);
    st.l = st.l * c1 + l2 * c2;
#pragma unroll
    for (int e = 0; e < 8; ++e) {
      const float a2 = __shfl_xor_sync(0xffffffffu, st.acc[e], off);
      st.acc[e] = st.acc[e] * c1 + a2 * c2;
    }
    st.m = mm;
  }
  if (sub == 0) {
    const float inv = 1.0f / st.l;
    float o[8];
#pragma unroll
    for (int e = 0; e < 8; ++e) o[e] = st.acc[e] * inv;
    row_store<8>(P.attn_o + (size_t)i * D + h * 64 + ch * 8, o);
  }
}

// While a QKV GEMM phase streams its weights, the compute warps pull the K/V rows their attention items of the same
// layer will read into L2 (the cached positions do not depend on this step): the attention phase then reads L2.
__device__ void kv_prefetch_l2(const MegaParams& P, int layer, int n_items, int cw, int lane) {
  const KvGeom& g = P.kv;
  for (int item = blockIdx.x + P.G * cw; item < n_items; item += P.G * MEGA_NCW) {
    const int i = item / P.H, h = item - i * P.H;
    const int slot = P.slot_ids[i];
    const int L = P.offsets[slot];
    const char* k0 = reinterpret_cast<const char*>(g.base) + ((((size_t)layer * g.n_slots + slot) * 2) * P.H + h) * (size_t)g.cap * 128;
    const char* v0 = k0 + (size_t)P.H * g.cap * 128;
    for (int p = lane; p < L; p += 32) {
      asm volatile("prefetch.global.L2 [%0];" ::"l"(k0 + (size_t)p * 128));
      asm volatile("prefetch.global.L2 [%0];" ::"l"(v0 + (size_t)p * 128));
    }
  }
}

// ------------------------------------------------------------------ compute warps: phase loop
__device__ __forceinline__ void phase_prefetch(const MegaParams& P, MegaPhase* sph, int p, int tid_c) {
  // asynchronous 16-byte copies of phase p's descriptor into its shared-memory slot (completed by cp.async.wait_all)
  constexpr int CHUNKS = (int)(sizeof(MegaPhase) / 16);
  if (p < P.n_phases && tid_c < CHUNKS) {
    const uint32_t dst = smem_u32(reinterpret_cast<const uint8_t*>(sph + (p & 1)) + tid_c * 16);
    const uint8_t* src = reinterpret_cast<const uint8_t*>(P.phases + p) + tid_c * 16;
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(dst), "l"(src) : "memory");
  }
}

__device__ void compute_role(const MegaParams& P, uint32_t tfull_a, uint32_t tempty_a, uint32_t tmem_base, float* red,
                             MegaPhase* sph, int warp, int lane) {
  const int cw = warp - 2;                       // 0..NCW-1
  const int tid_c = cw * 32 + lane;
  const int q = warp & 3, half = cw >> 2;        // TMEM lane quarter of this warp, column interleave among its peers
  uint32_t acc = 0;
  int done_upto = 0;
  if (P.trace && tid_c == 0) P.trace[(size_t)blockIdx.x * (P.n_phases + 1)] = globaltimer_ns();
  phase_prefetch(P, sph, 0, tid_c);
  asm volatile("cp.async.wait_all;" ::: "memory");
  cta_sync_phase();
  for (int p = 0; p < P.n_phases; ++p) {
    // the descriptor of this phase sits in shared memory (fetched during the previous phase); fetch the next one
    const MegaPhase& ph = sph[p & 1];
    phase_prefetch(P, sph, p + 1, tid_c);
    const int type = ph.type;
    const int n_items = ph.n_items >= 0 ? ph.n_items : -ph.n_items * P.n;   // negative: items per stream row
    if (type == MP_GEMM) {
      if (ph.flag) kv_prefetch_l2(P, ph.idx, P.H * P.n, cw, lane);
      for (int item = blockIdx.x; item < n_items; item += P.G, ++acc) {
        Item it;
        decode_item(&ph, item, it);
        const uint32_t buf = acc & 1u, use = acc >> 1;
        mbar_wait_bounded(tfull_a + 8u * buf, use & 1u, P, 0x40000u + (unsigned)p);
        tc_fence_after();
        if (q * 32 < P.rows_pad)
          gemm_epilogue(P, it, tmem_base + buf * 64u + ((uint32_t)(q * 32) << 16), q, half, lane);
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive_a(tempty_a + 8u * buf);
      }
    } else if (type == MP_ROWS && ph.op == MR_EMBED) {
      const int first = blockIdx.x + P.G * cw;
      if (first < n_items) {
        grid_wait(P, p, done_upto, lane, 0x50000u + (unsigned)p);
        for (int item = first; item < n_items; item += P.G * MEGA_NCW) row_embed(P, item, lane);
      }
    } else if (type == MP_ROWS) {
      if ((int)blockIdx.x < n_items) {
        if (cw == 0) grid_wait(P, p, done_upto, lane, 0x50000u + (unsigned)p);
        cta_sync_start();
        done_upto = p;
        for (int item = blockIdx.x; item < n_items; item += P.G) {
          switch (ph.op) {
            case MR_RES_LN: row_res_ln(P, ph, item, tid_c, cw, lane, red, false); break;
            case MR_FINAL: row_res_ln(P, ph, item, tid_c, cw, lane, red, true); break;
            case MR_SY: row_sy(P, ph, item, tid_c, lane); break;
            case MR_LNMOD: row_lnmod(P, ph, item, tid_c, cw, lane, red); break;
            default: row_commit(P, ph, item, tid_c); break;
          }
          if (item + P.G < n_items) cta_sync_start();   // reduction scratch reuse
        }
      }
    } else {
      // item, slot and position of this warp's first item do not depend on this step: fetch them before the wait
      int item = blockIdx.x + P.G * cw;
      if (item < n_items) {
        int slot = P.slot_ids[item / P.H];
        int pos = P.offsets[slot];
        grid_wait(P, p, done_upto, lane, 0x60000u + (unsigned)p);
        for (; item < n_items; item += P.G * MEGA_NCW) {
          attn_item(P, ph, item, slot, pos, lane);
          const int nx = item + P.G * MEGA_NCW;
          if (nx < n_items) { slot = P.slot_ids[nx / P.H]; pos = P.offsets[slot]; }
        }
      }
    }
    // phase p of this CTA is done: the CTA barrier orders every compute thread's writes before thread 0's
    // gpu-scope release (cumulative), which counts the CTA in
    asm volatile("cp.async.wait_all;" ::: "memory");
    cta_sync_phase();
    if (tid_c == 0) {
      red_release_u32(P.sync + p);
      if (P.trace) P.trace[(size_t)blockIdx.x * (P.n_phases + 1) + p + 1] = globaltimer_ns();
    }
  }
}

__global__ void __launch_bounds__(MEGA_THREADS, 2) flow_mega_kernel(const __grid_constant__ MegaParams P) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* tiles = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  const uint32_t stage_bytes = (uint32_t)P.a_bytes + MEGA_W_BYTES;
  float* red = reinterpret_cast<float*>(tiles + (size_t)P.stages * stage_bytes);
  __shared__ MegaPhase sph[2];
  __shared__ __align__(8) uint64_t full_bar[MEGA_MAX_STAGES];
  __shared__ __align__(8) uint64_t empty_bar[MEGA_MAX_STAGES];
  __shared__ __align__(8) uint64_t tfull_bar[2];
  __shared__ __align__(8) uint64_t tempty_bar[2];
  __shared__ uint32_t tmem_base_smem;

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (warp == 0 && lane == 0) {
    for (int s = 0; s < P.stages; ++s) { mbar_init(&full_bar[s], 1); mbar_init(&empty_bar[s], 1); }
    for (int b = 0; b < 2; ++b) { mbar_init(&tfull_bar[b], 1); mbar_init(&tempty_bar[b], MEGA_NCW); }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 1) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&tmem_base_smem)), "n"(MEGA_TMEM_COLS) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = tmem_base_smem;
  const uint32_t tiles_a = smem_u32(tiles), full_a = smem_u32(&full_bar[0]), empty_a = smem_u32(&empty_bar[0]);
  const uint32_t tfull_a = smem_u32(&tfull_bar[0]), tempty_a = smem_u32(&tempty_bar[0]);

  if (warp == 0) producer_role(P, tiles_a, full_a, empty_a, lane);
  else if (warp == 1) mma_role(P, tiles_a, full_a, empty_a, tfull_a, tempty_a, tmem_base, lane);
  else compute_role(P, tfull_a, tempty_a, tmem_base, red, sph, warp, lane);

  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "n"(MEGA_TMEM_COLS) : "memory");
  }
  // the last CTA out re-arms the phase counters for the next launch (every CTA has passed every poll)
  if (threadIdx.x == 0) {
    __threadfence();
    const unsigned old = atomicAdd(P.sync + MEGA_EXIT_WORD, 1u);
    if (old == (unsigned)P.G - 1u) {
      for (int p = 0; p < P.n_phases; ++p) P.sync[p] = 0u;
      P.sync[MEGA_EXIT_WORD] = 0u;
      __threadfence();
    }
  }
}

}  // namespace

size_t flow_mega_smem_bytes(int stages, int a_bytes) {
  return (size_t)stages * (a_bytes + MEGA_W_BYTES) + MEGA_SCRATCH + 1024;
}

cudaError_t flow_mega_launch(const MegaParams& P, cudaStream_t st) {
  if (P.stages < 2 || P.stages > MEGA_MAX_STAGES || P.n_phases > MEGA_MAX_PHASES) return cudaErrorInvalidValue;
  const size_t smem = flow_mega_smem_bytes(P.stages, P.a_bytes);
  static size_t attr_smem_dev[64] = {};
  int cur_dev = 0;
  cudaGetDevice(&cur_dev);
  size_t& attr_smem = attr_smem_dev[cur_dev & 63];
  if (smem > attr_smem) {   // static shared memory (barriers) counts against the 227 KB limit too
    cudaError_t e = cudaFuncSetAttribute(flow_mega_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
    attr_smem = smem;
  }
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = dim3(P.G, 1, 1);
  cfg.blockDim = dim3(MEGA_THREADS, 1, 1);
  cfg.dynamicSmemBytes = smem;
  cfg.stream = st;
  cfg.attrs = nullptr;
  cfg.numAttrs = 0;
  {  // the kernel spins on software grid barriers: every CTA must be resident at once
    static int max_ctas = -1;
    if (max_ctas < 0) {
      int per_sm = 0, sms = 0, dev = 0;
      cudaGetDevice(&dev);
      cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
      if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, flow_mega_kernel, (int)cfg.blockDim.x, cfg.dynamicSmemBytes) != cudaSuccess) per_sm = 0;
      max_ctas = per_sm * sms;
    }
    if ((int)cfg.gridDim.x > max_ctas) return cudaErrorCooperativeLaunchTooLarge;
  }
  return cudaLaunchKernelEx(&cfg, flow_mega_kernel, P);
}

}  // namespace ptts
