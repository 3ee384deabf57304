// Streaming attention kernels (transformer.py:135-158, rope.py:28-58) with per-stream
// offsets.  RoPE, the KV append and the softmax(QK^T)V read are fused: q/k/v of the new
// position(s) are read once from the packed projection, K/V are streamed once from HBM.
//
// Layout: one (stream, head) pair per CTA.  Each lane owns the interleaved RoPE pair
// (2*lane, 2*lane+1) of the 64-wide head, so a K or V row is one coalesced 256 B (fp32) /
// 128 B (bf16) warp load, the rotation needs no shuffles, and the score is a 5-step
// butterfly.  Warps split the positions and are merged with a log-sum-exp combine.
#include <algorithm>
#include <cstdlib>

#include "kernels.cuh"

namespace ptts {

PTTS_TL_ENABLE_FN(attention)

template <typename TK>
__device__ __forceinline__ TK* kv_row(const KvGeom& g, int layer, int slot, int kv, int h, int pos) {
  const int p = g.ring ? (pos & (g.cap - 1)) : pos;
  const long long idx = ((((long long)layer * g.n_slots + slot) * 2 + kv) * g.H + h) * g.cap + p;
  return reinterpret_cast<TK*>(g.base) + idx * g.D;
}

__device__ __forceinline__ float* kv_scale_row(const KvGeom& g, int layer, int slot, int kv, int h, int pos) {
  const long long idx = ((((long long)layer * g.n_slots + slot) * 2 + kv) * g.H + h) * g.cap + pos;
  return g.scale + idx;
}
template <typename TK> struct IsFp8 { static constexpr bool v = false; };
template <> struct IsFp8<fp8e4> { static constexpr bool v = true; };
// e4m3 row scale: amax / 448 (largest finite e4m3), 1 for an all-zero row
__device__ __forceinline__ float fp8_scale(float amax) { return amax > 0.0f ? amax * (1.0f / 448.0f) : 1.0f; }

// shared voice prefix of `slot`: rows [0, plen) live `delta` elements away (in the prefix owner's cache, same layout)
__device__ __forceinline__ long long pfx_delta(const KvGeom& g, int slot, int& plen) {
  plen = 0;
  if (!g.pfx_slot) return 0;
  const int ps = g.pfx_slot[slot];
  if (ps < 0) return 0;
  plen = g.pfx_len[slot];
  return (long long)(ps - slot) * 2 * g.H * g.cap * g.D;
}

__device__ __forceinline__ float2 rope_pair(float2 x, float pos, float freq) {
  // rope.py:42-50: angle = freq * position in fp32; (r, i) -> (r cos - i sin, r sin + i cos)
  float s, c;
  sincosf(freq * pos, &s, &c);
  return make_float2(x.x * c - x.y * s, x.x * s + x.y * c);
}

// online-softmax state of one warp for one query: running max m, sum l, accumulator pair
struct OnlineSm {
  float m, l;
  float2 acc;
};

// scores of 4 consecutive positions p..p+3 (count valid), lane-owned q pair, D == 64
template <typename TK>
__device__ __forceinline__ void attend4(OnlineSm& s, const TK* __restrict__ kbase,
                                        const TK* __restrict__ vbase, int idx0, int count,
                                        float2 q, float scale, int lane, int plen = 0, long long dsh = 0,
                                        const float* __restrict__ ksbase = nullptr, const float* __restrict__ vsbase = nullptr) {
  float2 kk[4], vv[4];
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    if (j < count) {
      const long long o = (long long)(idx0 + j) * 64 + 2 * lane + (idx0 + j < plen ? dsh : 0);
      kk[j] = load2(kbase + o);
      vv[j] = load2(vbase + o);
      if constexpr (IsFp8<TK>::v) {   // FP8 cache: rows carry their scale
        const long long so = (long long)(idx0 + j) + (idx0 + j < plen ? dsh / 64 : 0);
        const float ks = ksbase[so], vs = vsbase[so];
        kk[j].x *= ks; kk[j].y *= ks; vv[j].x *= vs; vv[j].y *= vs;
      }
    } else {
      kk[j] = make_float2(0.f, 0.f);
      vv[j] = make_float2(0.f, 0.f);
    }
  }
  float sc[4];
#pragma unroll
  for (int j = 0; j < 4; ++j) sc[j] = q.x * kk[j].x + q.y * kk[j].y;
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
#pragma unroll
    for (int j = 0; j < 4; ++j) sc[j] += __shfl_xor_sync(0xffffffffu, sc[j], o);
  }
  float mx = s.m;
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    sc[j] = (j < count) ? sc[j] * scale : -INFINITY;
    mx = fmaxf(mx, sc[j]);
  }
  const float corr = expf(s.m - mx);  // s.m == -inf on first use -> 0
  s.l *= corr;
  s.acc.x *= corr;
  s.acc.y *= corr;
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    const float p = expf(sc[j] - mx);  // -inf -> 0
    s.l += p;
    s.acc.x = fmaf(p, vv[j].x, s.acc.x);
    s.acc.y = fmaf(p, vv[j].y, s.acc.y);
  }
  s.m = mx;
}

// =====================================================================================
// decode: one new position per stream.  grid (H, n), 128 threads.
// A lane owns a 16-byte slice of a K / V row (8 bf16 or 4 fp32 dims) and a 32-lane load covers
// RPL = 4 (bf16) or 2 (fp32) consecutive positions, so K/V stream in as fully coalesced 512-byte
// requests.  Every lane group keeps its own online-softmax state for the positions it sees (no
// per-iteration cross-lane max / sum); U loads of K and of V are in flight per warp before any is
// consumed.  Groups are merged once at the end (shuffles, then shared memory across warps).
// =====================================================================================
template <typename TK> struct KvVec;
template <> struct KvVec<bf16> {
  static constexpr int EPL = 8;   // elements per lane
  typedef uint4 Raw;
  __device__ static Raw load_raw(const bf16* p) { return *reinterpret_cast<const uint4*>(p); }
  __device__ static Raw zero() { return make_uint4(0u, 0u, 0u, 0u); }
  __device__ static void unpack(const Raw& r, float* f) {
    const uint32_t w[4] = {r.x, r.y, r.z, r.w};
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const float2 t = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&w[i]));
      f[2 * i] = t.x; f[2 * i + 1] = t.y;
    }
  }
  __device__ static void load(const bf16* p, float* f) {
    const uint4 r = *reinterpret_cast<const uint4*>(p);
    const uint32_t w[4] = {r.x, r.y, r.z, r.w};
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const float2 t = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&w[i]));
      f[2 * i] = t.x; f[2 * i + 1] = t.y;
    }
  }
  __device__ static void store(bf16* p, const float* f) {
    uint32_t w[4];
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const __nv_bfloat162 t = __floats2bfloat162_rn(f[2 * i], f[2 * i + 1]);
      w[i] = *reinterpret_cast<const uint32_t*>(&t);
    }
    *reinterpret_cast<uint4*>(p) = make_uint4(w[0], w[1], w[2], w[3]);
  }
};
template <> struct KvVec<fp8e4> {
  static constexpr int EPL = 16;  // 16 bytes per lane: a 64-wide row is 4 lanes, a warp load covers 8 positions
  typedef uint4 Raw;
  __device__ static Raw load_raw(const fp8e4* p) { return *reinterpret_cast<const uint4*>(p); }
  __device__ static Raw zero() { return make_uint4(0u, 0u, 0u, 0u); }
  __device__ static void unpack(const Raw& r, float* f) {
    const uint32_t w[4] = {r.x, r.y, r.z, r.w};
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const float2 lo = load2(reinterpret_cast<const fp8e4*>(&w[i]));
      const float2 hi = load2(reinterpret_cast<const fp8e4*>(&w[i]) + 2);
      f[4 * i] = lo.x; f[4 * i + 1] = lo.y; f[4 * i + 2] = hi.x; f[4 * i + 3] = hi.y;
    }
  }
  __device__ static void store(fp8e4* p, const float* f) {   // values already divided by the row scale
    uint32_t w[4];
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const uint32_t lo = __nv_cvt_float2_to_fp8x2(make_float2(f[4 * i], f[4 * i + 1]), __NV_SATFINITE, __NV_E4M3);
      const uint32_t hi = __nv_cvt_float2_to_fp8x2(make_float2(f[4 * i + 2], f[4 * i + 3]), __NV_SATFINITE, __NV_E4M3);
      w[i] = lo | (hi << 16);
    }
    *reinterpret_cast<uint4*>(p) = make_uint4(w[0], w[1], w[2], w[3]);
  }
};
template <> struct KvVec<float> {
  static constexpr int EPL = 4;
  typedef float4 Raw;
  __device__ static Raw load_raw(const float* p) { return *reinterpret_cast<const float4*>(p); }
  __device__ static Raw zero() { return make_float4(0.f, 0.f, 0.f, 0.f); }
  __device__ static void unpack(const Raw& r, float* f) { f[0] = r.x; f[1] = r.y; f[2] = r.z; f[3] = r.w; }
  __device__ static void load(const float* p, float* f) {
    const float4 r = *reinterpret_cast<const float4*>(p);
    f[0] = r.x; f[1] = r.y; f[2] = r.z; f[3] = r.w;
  }
  __device__ static void store(float* p, const float* f) { *reinterpret_cast<float4*>(p) = make_float4(f[0], f[1], f[2], f[3]); }
};

template <typename TK, typename TO, int U, bool SH = false>   // U: 16-byte loads of K and of V in flight per lane (rows stay packed in registers); SH: some stream reads a shared voice prefix
__global__ void __launch_bounds__(128)   // bf16: U = 6 -> 72 registers, 7 CTAs per SM; U = 4 -> 56; U = 8 -> 80
attn_decode_kernel(const float* __restrict__ qkv, const float* __restrict__ freqs,
                   const int* __restrict__ offsets, const int* __restrict__ slot_ids, KvGeom g,
                   int layer, TO* __restrict__ out) {
  pdl_launch_dependents();
  constexpr int NW = 4;
  constexpr int EPL = KvVec<TK>::EPL;      // dims per lane
  constexpr int LPR = 64 / EPL;            // lanes per row: 8 / 16
  constexpr int RPL = 32 / LPR;            // rows per warp load: 4 / 2
  __shared__ float sm_m[NW][32], sm_l[NW][32];
  __shared__ float sm_acc[NW][32][EPL];
  const int h = blockIdx.x, i = blockIdx.y;
  const int slot = slot_ids[i];
  const int pos = offsets[slot];           // advanced only by the last kernel of the step: stable here
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  const int sub = lane / LPR, ch = lane % LPR;   // row within a load, 16-byte chunk of the row
  const int d_model = g.H * 64;
  int plen = 0;
  const long long dsh = SH ? pfx_delta(g, slot, plen) : 0;
  if (g.prefetch) {
    // The cached positions [0, pos) were written by earlier steps, not by the kernel this one waits for: pull
    // them into L2 now, while the QKV projection (latency-bound, HBM nearly idle) is still running.
    const char* kb = reinterpret_cast<const char*>(kv_row<TK>(g, layer, slot, 0, h, 0));
    const char* vb = reinterpret_cast<const char*>(kv_row<TK>(g, layer, slot, 1, h, 0));
    const int bytes = pos * 64 * (int)sizeof(TK), pbytes = plen * 64 * (int)sizeof(TK);
    for (int off = threadIdx.x * 128; off < bytes; off += 128 * 128) {
      const long long sh = (SH && off < pbytes) ? dsh * (long long)sizeof(TK) : 0;
      asm volatile("prefetch.global.L2 [%0];" ::"l"(kb + off + sh));
      asm volatile("prefetch.global.L2 [%0];" ::"l"(vb + off + sh));
    }
  }
  const unsigned tl = tl_begin(2);
  pdl_wait();
  tl_mark(tl, 2);
  const float* row = qkv + (long long)i * 3 * d_model + h * 64 + ch * EPL;

  // rotate this lane's slice of q (interleaved pairs, rope.py:36-58); lanes with sub == 0 of
  // warp 0 append the rotated k and v
  float q[EPL];
  {
    float kx[EPL], vx[EPL];
    if constexpr (IsFp8<TK>::v) {
#pragma unroll
      for (int e = 0; e < EPL; ++e) { kx[e] = 0.0f; vx[e] = 0.0f; }
    }
#pragma unroll
    for (int e = 0; e < EPL; e += 2) {
      const float fr = freqs[(ch * EPL + e) >> 1];
      const float2 qq = rope_pair(load2(row + e), (float)pos, fr);
      q[e] = qq.x * 0.125f; q[e + 1] = qq.y * 0.125f;   // fold the 1/sqrt(64) softmax scale into q
      if (w == 0 && sub == 0) {
        const float2 kk = rope_pair(load2(row + d_model + e), (float)pos, fr);
        const float2 vv = load2(row + 2 * d_model + e);
        kx[e] = kk.x; kx[e + 1] = kk.y; vx[e] = vv.x; vx[e + 1] = vv.y;
      }
    }
    if constexpr (IsFp8<TK>::v) {
      // row scale = amax / 448 over the 64 dims of the row (the LPR lanes of lane group 0 hold them); whole warp 0 runs
      // the shuffles, lane group 0 stores
      if (w == 0) {
        float ka = 0.0f, va = 0.0f;
#pragma unroll
        for (int e = 0; e < EPL; ++e) { ka = fmaxf(ka, fabsf(kx[e])); va = fmaxf(va, fabsf(vx[e])); }
#pragma unroll
        for (int o = 1; o < LPR; o <<= 1) {
          ka = fmaxf(ka, __shfl_xor_sync(0xffffffffu, ka, o));
          va = fmaxf(va, __shfl_xor_sync(0xffffffffu, va, o));
        }
        if (sub == 0) {
          const float ks = fp8_scale(ka), vs = fp8_scale(va), kinv = 1.0f / ks, vinv = 1.0f / vs;
#pragma unroll
          for (int e = 0; e < EPL; ++e) { kx[e] *= kinv; vx[e] *= vinv; }
          KvVec<TK>::store(kv_row<TK>(g, layer, slot, 0, h, pos) + ch * EPL, kx);
          KvVec<TK>::store(kv_row<TK>(g, layer, slot, 1, h, pos) + ch * EPL, vx);
          if (ch == 0) { *kv_scale_row(g, layer, slot, 0, h, pos) = ks; *kv_scale_row(g, layer, slot, 1, h, pos) = vs; }
        }
      }
    } else if (w == 0 && sub == 0) {
      KvVec<TK>::store(kv_row<TK>(g, layer, slot, 0, h, pos) + ch * EPL, kx);
      KvVec<TK>::store(kv_row<TK>(g, layer, slot, 1, h, pos) + ch * EPL, vx);
    }
  }
  __syncthreads();  // appended row visible to the block

  const TK* kbase = kv_row<TK>(g, layer, slot, 0, h, 0) + ch * EPL;
  const TK* vbase = kv_row<TK>(g, layer, slot, 1, h, 0) + ch * EPL;
  const float* ksbase = IsFp8<TK>::v ? kv_scale_row(g, layer, slot, 0, h, 0) : nullptr;
  const float* vsbase = IsFp8<TK>::v ? kv_scale_row(g, layer, slot, 1, h, 0) : nullptr;
  const int L = pos + 1;
  float m = -INFINITY, l = 0.0f, acc[EPL];
#pragma unroll
  for (int e = 0; e < EPL; ++e) acc[e] = 0.0f;
  constexpr int STEP = NW * U * RPL;   // positions per block iteration
  for (int p0 = w * U * RPL; p0 < L; p0 += STEP) {
    typename KvVec<TK>::Raw kr[U], vr[U];   // rows stay packed until they are used: U loads of K and of V in flight
    float ksc[IsFp8<TK>::v ? U : 1], vsc[IsFp8<TK>::v ? U : 1];
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const int p = p0 + u * RPL + sub;
      if (p < L) {
        const long long o = (long long)p * 64 + ((SH && p < plen) ? dsh : 0);
        kr[u] = KvVec<TK>::load_raw(kbase + o);
        vr[u] = KvVec<TK>::load_raw(vbase + o);
        if constexpr (IsFp8<TK>::v) {
          const long long so = (long long)p + ((SH && p < plen) ? dsh / 64 : 0);
          ksc[u] = ksbase[so]; vsc[u] = vsbase[so];
        }
      } else {
        kr[u] = KvVec<TK>::zero(); vr[u] = KvVec<TK>::zero();
        if constexpr (IsFp8<TK>::v) { ksc[u] = 0.0f; vsc[u] = 0.0f; }
      }
    }
    float sc[U];
#pragma unroll
    for (int u = 0; u < U; ++u) {
      float kf[EPL];
      KvVec<TK>::unpack(kr[u], kf);
      float a = 0.0f;
#pragma unroll
      for (int e = 0; e < EPL; ++e) a = fmaf(q[e], kf[e], a);
      sc[u] = a;
    }
#pragma unroll
    for (int o = 1; o < LPR; o <<= 1) {
#pragma unroll
      for (int u = 0; u < U; ++u) sc[u] += __shfl_xor_sync(0xffffffffu, sc[u], o);
    }
    float mx = m;
#pragma unroll
    for (int u = 0; u < U; ++u) {
      if constexpr (IsFp8<TK>::v) sc[u] *= ksc[u];
      if (p0 + u * RPL + sub >= L) sc[u] = -INFINITY;
      mx = fmaxf(mx, sc[u]);
    }
    if (mx > -INFINITY) {   // this lane group has seen at least one position
      const float corr = expf(m - mx);
      l *= corr;
#pragma unroll
      for (int e = 0; e < EPL; ++e) acc[e] *= corr;
#pragma unroll
      for (int u = 0; u < U; ++u) {
        const float pr = expf(sc[u] - mx);
        l += pr;
        float vf[EPL];
        KvVec<TK>::unpack(vr[u], vf);
        const float pv = IsFp8<TK>::v ? pr * vsc[IsFp8<TK>::v ? u : 0] : pr;
#pragma unroll
        for (int e = 0; e < EPL; ++e) acc[e] = fmaf(pv, vf[e], acc[e]);
      }
      m = mx;
    }
  }
  sm_m[w][lane] = m;
  sm_l[w][lane] = l;
#pragma unroll
  for (int e = 0; e < EPL; ++e) sm_acc[w][lane][e] = acc[e];
  __syncthreads();
  if (w == 0 && sub == 0) {
    // merge the NW * RPL partial states that share this chunk of dims
    float mm = -INFINITY;
    for (int j = 0; j < NW; ++j)
      for (int r = 0; r < RPL; ++r) mm = fmaxf(mm, sm_m[j][r * LPR + ch]);
    float ll = 0.0f, a[EPL];
#pragma unroll
    for (int e = 0; e < EPL; ++e) a[e] = 0.0f;
    for (int j = 0; j < NW; ++j)
      for (int r = 0; r < RPL; ++r) {
        const int src = r * LPR + ch;
        const float c = expf(sm_m[j][src] - mm);   // empty partials hold m = -inf -> 0
        ll += sm_l[j][src] * c;
#pragma unroll
        for (int e = 0; e < EPL; ++e) a[e] = fmaf(sm_acc[j][src][e], c, a[e]);
      }
    const float inv = 1.0f / ll;
    TO* o = out + (long long)i * d_model + h * 64 + ch * EPL;
#pragma unroll
    for (int e = 0; e < EPL; e += 2) store2(o + e, make_float2(a[e] * inv, a[e + 1] * inv));
  }
  tl_mark(tl, 3);
  if (tl_buf_dev && threadIdx.x == 0) atomicMax(tl_buf_dev + 1, tl_now());   // latest exit of any CTA of any launch
}

// =====================================================================================
// decode, bulk-copy form (default): one (stream, head) per WARP.  The warp streams the cached K and V rows of
// its pair — two contiguous runs of pos * 128 bytes (bf16) — through a private shared-memory ring with
// cp.async.bulk (4 KB of K + 4 KB of V per chunk, completion on an mbarrier), so the whole cache read is a
// handful of TMA requests per warp instead of hundreds of per-lane loads, and it is issued BEFORE the kernel
// waits for the QKV projection: cached rows do not depend on it (offsets advance only in the last kernel of a
// step, and steps are separated by a graph / copy boundary), so under programmatic dependent launch the KV
// stream overlaps the latency-bound GEMM in front of it.  The new position is folded in from registers (rounded
// through the cache type, as if re-read) and appended to the cache.  No block-level synchronisation at all.
// =====================================================================================
constexpr int AD_CHUNK = 4096;   // bytes of K (and of V) per ring slot
constexpr int AD_SLOTS = 2;

__device__ __forceinline__ void bulk_g2s(uint32_t dst, const void* src, uint32_t bytes, uint32_t bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
               ::"r"(dst), "l"(src), "r"(bytes), "r"(bar) : "memory");
}
__device__ __forceinline__ void ad_wait(uint32_t bar, uint32_t parity) {
  uint32_t ok = 0;
  for (uint32_t spin = 0; spin < (1u << 22) && !ok; ++spin)
    asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                 : "=r"(ok) : "r"(bar), "r"(parity) : "memory");
  if (!ok) __trap();
}

template <typename TK, typename TO, int W>
__global__ void __launch_bounds__(W * 32)
attn_decode_bulk_kernel(const float* __restrict__ qkv, const float* __restrict__ freqs,
                        const int* __restrict__ offsets, const int* __restrict__ slot_ids, KvGeom g,
                        int layer, TO* __restrict__ out, int n_tasks) {
  constexpr int EPL = KvVec<TK>::EPL;      // dims per lane
  constexpr int LPR = 64 / EPL;            // lanes per row: 8 / 16
  constexpr int RPL = 32 / LPR;            // rows per warp load: 4 / 2
  constexpr int U = 4;
  constexpr int ROWB = 64 * (int)sizeof(TK);
  constexpr int CR = AD_CHUNK / ROWB;      // rows per chunk: 32 / 16
  typedef typename KvVec<TK>::Raw Raw;
  extern __shared__ __align__(128) uint8_t ad_smem[];   // [W][AD_SLOTS][K | V][AD_CHUNK]
  __shared__ __align__(8) uint64_t bars[W][AD_SLOTS];
  pdl_launch_dependents();
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  const int task = blockIdx.x * W + w;
  if (task >= n_tasks) return;             // whole warp; nothing below synchronises across warps
  const int i = task / g.H, h = task - i * g.H;
  const int slot = slot_ids[i];
  const int pos = offsets[slot];
  const int sub = lane / LPR, ch = lane % LPR;
  const int d_model = g.H * 64;
  const uint32_t ring = (uint32_t)__cvta_generic_to_shared(ad_smem) + (uint32_t)w * AD_SLOTS * 2 * AD_CHUNK;
  const uint32_t bar0 = (uint32_t)__cvta_generic_to_shared(&bars[w][0]);
  const char* kg = reinterpret_cast<const char*>(kv_row<TK>(g, layer, slot, 0, h, 0));
  const char* vg = reinterpret_cast<const char*>(kv_row<TK>(g, layer, slot, 1, h, 0));
  const int n_chunks = (pos + CR - 1) / CR;
  auto issue = [&](int c) {                // lane 0: chunk c -> ring slot c % AD_SLOTS
    const int s = c % AD_SLOTS;
    const uint32_t bytes = (uint32_t)min(CR, pos - c * CR) * ROWB;
    const uint32_t dst = ring + (uint32_t)s * 2 * AD_CHUNK, bar = bar0 + 8u * s;
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(2 * bytes) : "memory");
    bulk_g2s(dst, kg + (size_t)c * AD_CHUNK, bytes, bar);
    bulk_g2s(dst + AD_CHUNK, vg + (size_t)c * AD_CHUNK, bytes, bar);
  };
  if (lane == 0) {
    for (int s = 0; s < AD_SLOTS; ++s) asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(bar0 + 8u * s));
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    for (int c = 0; c < AD_SLOTS && c < n_chunks; ++c) issue(c);
  }
  __syncwarp();
  pdl_wait();

  // rotate this lane's slice of q and of the new k (interleaved pairs, rope.py:36-58); lane group 0 appends
  const float* row = qkv + (long long)i * 3 * d_model + h * 64 + ch * EPL;
  float q[EPL], kn[EPL], vn[EPL];
#pragma unroll
  for (int e = 0; e < EPL; e += 2) {
    const float fr = freqs[(ch * EPL + e) >> 1];
    const float2 qq = rope_pair(load2(row + e), (float)pos, fr);
    q[e] = qq.x * 0.125f; q[e + 1] = qq.y * 0.125f;   // fold the 1/sqrt(64) softmax scale into q
    const float2 kk = rope_pair(load2(row + d_model + e), (float)pos, fr);
    const float2 vv = load2(row + 2 * d_model + e);
    kn[e] = to_f(from_f<TK>(kk.x)); kn[e + 1] = to_f(from_f<TK>(kk.y));   // as stored in the cache
    vn[e] = to_f(from_f<TK>(vv.x)); vn[e + 1] = to_f(from_f<TK>(vv.y));
  }
  if (sub == 0) {
    KvVec<TK>::store(kv_row<TK>(g, layer, slot, 0, h, pos) + ch * EPL, kn);
    KvVec<TK>::store(kv_row<TK>(g, layer, slot, 1, h, pos) + ch * EPL, vn);
  }

  float m = -INFINITY, l = 0.0f, acc[EPL];
#pragma unroll
  for (int e = 0; e < EPL; ++e) acc[e] = 0.0f;
  for (int c = 0; c < n_chunks; ++c) {
    const int s = c % AD_SLOTS;
    ad_wait(bar0 + 8u * s, (uint32_t)(c / AD_SLOTS) & 1u);
    const int rows = min(CR, pos - c * CR);
    const uint8_t* kb = ad_smem + ((size_t)w * AD_SLOTS + s) * 2 * AD_CHUNK + ch * 16;
    const uint8_t* vb = kb + AD_CHUNK;
#pragma unroll
    for (int it = 0; it < CR / (U * RPL); ++it) {
      if (it * U * RPL >= rows) break;     // uniform
      float sc[U];
      Raw vr[U];
#pragma unroll
      for (int u = 0; u < U; ++u) {
        const int r = it * U * RPL + u * RPL + sub;
        float a = 0.0f;
        if (r < rows) {
          float kf[EPL];
          KvVec<TK>::unpack(*reinterpret_cast<const Raw*>(kb + r * ROWB), kf);
          vr[u] = *reinterpret_cast<const Raw*>(vb + r * ROWB);
#pragma unroll
          for (int e = 0; e < EPL; ++e) a = fmaf(q[e], kf[e], a);
        } else {
          vr[u] = KvVec<TK>::zero();
        }
        sc[u] = a;
      }
#pragma unroll
      for (int o = 1; o < LPR; o <<= 1) {
#pragma unroll
        for (int u = 0; u < U; ++u) sc[u] += __shfl_xor_sync(0xffffffffu, sc[u], o);
      }
      float mx = m;
#pragma unroll
      for (int u = 0; u < U; ++u) {
        if (it * U * RPL + u * RPL + sub >= rows) sc[u] = -INFINITY;
        mx = fmaxf(mx, sc[u]);
      }
      if (mx > -INFINITY) {
        const float corr = expf(m - mx);
        l *= corr;
#pragma unroll
        for (int e = 0; e < EPL; ++e) acc[e] *= corr;
#pragma unroll
        for (int u = 0; u < U; ++u) {
          const float pr = expf(sc[u] - mx);
          l += pr;
          float vf[EPL];
          KvVec<TK>::unpack(vr[u], vf);
#pragma unroll
          for (int e = 0; e < EPL; ++e) acc[e] = fmaf(pr, vf[e], acc[e]);
        }
        m = mx;
      }
    }
    __syncwarp();                          // every lane is done with this slot
    if (lane == 0 && c + AD_SLOTS < n_chunks) {
      asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
      issue(c + AD_SLOTS);
    }
  }
  {  // the new position: lane group 0 takes it into its state
    float a = 0.0f;
#pragma unroll
    for (int e = 0; e < EPL; ++e) a = fmaf(q[e], kn[e], a);
#pragma unroll
    for (int o = 1; o < LPR; o <<= 1) a += __shfl_xor_sync(0xffffffffu, a, o);
    if (sub == 0) {
      const float mx = fmaxf(m, a);
      const float corr = expf(m - mx), pr = expf(a - mx);
      l = l * corr + pr;
#pragma unroll
      for (int e = 0; e < EPL; ++e) acc[e] = fmaf(pr, vn[e], acc[e] * corr);
      m = mx;
    }
  }
  // merge the RPL lane groups (they hold the same dims for different positions)
#pragma unroll
  for (int o = LPR; o < 32; o <<= 1) {
    const float m2 = __shfl_xor_sync(0xffffffffu, m, o), l2 = __shfl_xor_sync(0xffffffffu, l, o);
    const float mx = fmaxf(m, m2);
    const float c1 = (m > -INFINITY) ? expf(m - mx) : 0.0f, c2 = (m2 > -INFINITY) ? expf(m2 - mx) : 0.0f;
    l = l * c1 + l2 * c2;
#pragma unroll
    for (int e = 0; e < EPL; ++e) {
      const float a2 = __shfl_xor_sync(0xffffffffu, acc[e], o);
      acc[e] = acc[e] * c1 + a2 * c2;
    }
    m = mx;
  }
  if (sub == 0) {
    const float inv = 1.0f / l;
    TO* o = out + (long long)i * d_model + h * 64 + ch * EPL;
#pragma unroll
    for (int e = 0; e < EPL; e += 2) store2(o + e, make_float2(acc[e] * inv, acc[e + 1] * inv));
  }
}

template <typename TK, typename TO, int W>
static cudaError_t launch_attn_decode_bulk_w(const float* qkv, const float* rope_freqs, const int* offsets, const int* slot_ids,
                                             KvGeom kv, int layer, TO* out, int n, cudaStream_t st) {
  const int tasks = n * kv.H;
  const size_t smem = (size_t)W * AD_SLOTS * 2 * AD_CHUNK;
  auto kern = attn_decode_bulk_kernel<TK, TO, W>;
  static DeviceOnce attr_done;
  if (attr_done.first()) {
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
  }
  launch_k(kern, dim3((tasks + W - 1) / W), dim3(W * 32), smem, st, qkv, rope_freqs, offsets, slot_ids, kv, layer, out, tasks);
  return cudaGetLastError();
}
template <typename TK, typename TO>
static cudaError_t launch_attn_decode_bulk(const float* qkv, const float* rope_freqs, const int* offsets, const int* slot_ids,
                                           KvGeom kv, int layer, TO* out, int n, cudaStream_t st) {
  // warps (= (stream, head) pairs) per CTA: 7 puts the 1024 pairs of 64 streams on 147 CTAs, one per SM
  static const int w = [] { const char* e = getenv("PTTS_ATTN_W"); return e ? atoi(e) : 7; }();
  if (w == 2) return launch_attn_decode_bulk_w<TK, TO, 2>(qkv, rope_freqs, offsets, slot_ids, kv, layer, out, n, st);
  if (w == 4) return launch_attn_decode_bulk_w<TK, TO, 4>(qkv, rope_freqs, offsets, slot_ids, kv, layer, out, n, st);
  return launch_attn_decode_bulk_w<TK, TO, 7>(qkv, rope_freqs, offsets, slot_ids, kv, layer, out, n, st);
}

cudaError_t launch_attn_decode(const float* qkv, const float* rope_freqs, const int* offsets,
                               const int* slot_ids, KvGeom kv, int layer, void* out, int out_type,
                               int n, cudaStream_t st) {
  if (kv.D != 64) return cudaErrorInvalidValue;
  if (kv.type == 2) {   // FP8 cache: per-lane form (16 dims per lane), with or without a shared prefix
    if (out_type != 1 || !kv.scale) return cudaErrorInvalidValue;
    dim3 grid(kv.H, n);
    if (kv.pfx_slot) launch_k(attn_decode_kernel<fp8e4, bf16, 4, true>, grid, 128, 0, st, qkv, rope_freqs, offsets, slot_ids, kv, layer, (bf16*)out);
    else launch_k(attn_decode_kernel<fp8e4, bf16, 4, false>, grid, 128, 0, st, qkv, rope_freqs, offsets, slot_ids, kv, layer, (bf16*)out);
    return cudaGetLastError();
  }
  if (kv.pfx_slot) {   // some stream of the batch reads a shared voice prefix: per-lane form with the row select
    dim3 grid(kv.H, n);
    if (kv.type == 0 && out_type == 0)
      launch_k(attn_decode_kernel<float, float, 6, true>, grid, 128, 0, st, qkv, rope_freqs, offsets, slot_ids, kv, layer, (float*)out);
    else if (kv.type == 1 && out_type == 1)
      launch_k(attn_decode_kernel<bf16, bf16, 6, true>, grid, 128, 0, st, qkv, rope_freqs, offsets, slot_ids, kv, layer, (bf16*)out);
    else if (kv.type == 0 && out_type == 1)
      launch_k(attn_decode_kernel<float, bf16, 6, true>, grid, 128, 0, st, qkv, rope_freqs, offsets, slot_ids, kv, layer, (bf16*)out);
    else
      return cudaErrorInvalidValue;
    return cudaGetLastError();
  }
  if (kv.prefetch == 2 && !kv.ring) {
    if (kv.type == 0 && out_type == 0) return launch_attn_decode_bulk<float, float>(qkv, rope_freqs, offsets, slot_ids, kv, layer, (float*)out, n, st);
    if (kv.type == 1 && out_type == 1) return launch_attn_decode_bulk<bf16, bf16>(qkv, rope_freqs, offsets, slot_ids, kv, layer, (bf16*)out, n, st);
    if (kv.type == 0 && out_type == 1) return launch_attn_decode_bulk<float, bf16>(qkv, rope_freqs, offsets, slot_ids, kv, layer, (bf16*)out, n, st);
    return cudaErrorInvalidValue;
  }
  dim3 grid(kv.H, n);
  static const int u_sel = [] { const char* e = getenv("PTTS_ATTN_U"); return e ? atoi(e) : 6; }();
#define ATTN_CASE(UV)                                                                                                              \
  if (kv.type == 0 && out_type == 0)                                                                                               \
    launch_k(attn_decode_kernel<float, float, UV>, grid, 128, 0, st, qkv, rope_freqs, offsets, slot_ids, kv, layer, (float*)out);  \
  else if (kv.type == 1 && out_type == 1)                                                                                          \
    launch_k(attn_decode_kernel<bf16, bf16, UV>, grid, 128, 0, st, qkv, rope_freqs, offsets, slot_ids, kv, layer, (bf16*)out);     \
  else if (kv.type == 0 && out_type == 1)                                                                                          \
    launch_k(attn_decode_kernel<float, bf16, UV>, grid, 128, 0, st, qkv, rope_freqs, offsets, slot_ids, kv, layer, (bf16*)out);    \
  else                                                                                                                             \
    return cudaErrorInvalidValue;
  if (u_sel == 4) { ATTN_CASE(4) } else if (u_sel == 8) { ATTN_CASE(8) } else { ATTN_CASE(6) }
#undef ATTN_CASE
  return cudaGetLastError();
}

}  // namespace ptts
#include "attn_window_mma.cuh"
#include "attn_prefill_mma.cuh"
namespace ptts {

// =====================================================================================
// chunk: len[i] new positions per stream (prefill, Mimi's 16 tokens per frame).
// grid (H, n), 256 threads.  Phase 1 rotates q in place and appends rotated k / v;
// phase 2: one warp per query, causal + optional window (transformer.py:22-29).
// =====================================================================================
template <typename TK, typename TO>
__global__ void __launch_bounds__(256)
attn_chunk_kernel(float* __restrict__ qkv, const float* __restrict__ freqs,
                  const int* __restrict__ offsets, const int* __restrict__ slot_ids,
                  const int* __restrict__ row_start, const int* __restrict__ lens, KvGeom g, int layer,
                  int context, TO* __restrict__ out, int mode) {
  // mode 0: rotate / append, block barrier, attend (one CTA per (head, stream)); mode 1: rotate / append only;
  // mode 2: attend only, query t = blockIdx.z * 8 + warp (launched after a mode-1 pass: every query of a long
  // chunk gets its own warp somewhere on the machine instead of queueing behind 8 warps of one CTA)
  pdl_launch_dependents();
  pdl_wait();
  constexpr int NW = 8;
  const int h = blockIdx.x, i = blockIdx.y;
  const int slot = slot_ids[i];
  const int off = offsets[slot];
  const int r0 = row_start[i], T = lens[i];
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  const int d_model = g.H * 64;
  const float fr = freqs[lane];

  if (mode != 2)
  for (int t = w; t < T; t += NW) {
    float* row = qkv + (long long)(r0 + t) * 3 * d_model + h * 64 + 2 * lane;
    const float pos = (float)(off + t);
    store2(row, rope_pair(load2(row), pos, fr));
    float2 k = rope_pair(load2(row + d_model), pos, fr);
    float2 v = load2(row + 2 * d_model);
    if constexpr (IsFp8<TK>::v) {
      const float ks = fp8_scale(warp_max(fmaxf(fabsf(k.x), fabsf(k.y)))), vs = fp8_scale(warp_max(fmaxf(fabsf(v.x), fabsf(v.y))));
      k.x /= ks; k.y /= ks; v.x /= vs; v.y /= vs;
      if (lane == 0) { *kv_scale_row(g, layer, slot, 0, h, off + t) = ks; *kv_scale_row(g, layer, slot, 1, h, off + t) = vs; }
    }
    store2(kv_row<TK>(g, layer, slot, 0, h, off + t) + 2 * lane, k);
    store2(kv_row<TK>(g, layer, slot, 1, h, off + t) + 2 * lane, v);
  }
  if (mode == 1) return;
  __syncthreads();

  const TK* kbase = kv_row<TK>(g, layer, slot, 0, h, 0);
  const TK* vbase = kv_row<TK>(g, layer, slot, 1, h, 0);
  int plen = 0;
  const long long dsh = pfx_delta(g, slot, plen);   // shared voice prefix (never with ring storage)
  const float* ksbase = IsFp8<TK>::v ? kv_scale_row(g, layer, slot, 0, h, 0) : nullptr;
  const float* vsbase = IsFp8<TK>::v ? kv_scale_row(g, layer, slot, 1, h, 0) : nullptr;
  const int t_first = mode == 2 ? (int)blockIdx.z * NW + w : w;
  const int t_step = mode == 2 ? T : NW;        // mode 2: exactly one query per warp
  for (int t = t_first; t < T; t += t_step) {
    const float* row = qkv + (long long)(r0 + t) * 3 * d_model + h * 64 + 2 * lane;
    const float2 q = load2(row);
    const int pq = off + t;
    int lo = pq - context + 1;
    if (lo < 0) lo = 0;
    OnlineSm s{-INFINITY, 0.0f, make_float2(0.f, 0.f)};
    for (int p0 = lo; p0 <= pq; p0 += 4) {
      int cnt = min(4, pq + 1 - p0);
      if (g.ring) {
        // ring storage: the 4 positions may wrap; handle one at a time when they do
        const int i0 = p0 & (g.cap - 1);
        if (i0 + cnt > g.cap) {
          for (int j = 0; j < cnt; ++j) attend4<TK>(s, kbase, vbase, (p0 + j) & (g.cap - 1), 1, q, 0.125f, lane);
        } else {
          attend4<TK>(s, kbase, vbase, i0, cnt, q, 0.125f, lane);
        }
      } else {
        attend4<TK>(s, kbase, vbase, p0, cnt, q, 0.125f, lane, plen, dsh, ksbase, vsbase);
      }
    }
    const float inv = 1.0f / s.l;
    store2(out + (long long)(r0 + t) * d_model + h * 64 + 2 * lane, make_float2(s.acc.x * inv, s.acc.y * inv));
  }
}

cudaError_t launch_attn_chunk(float* qkv, const float* rope_freqs, const int* offsets,
                              const int* slot_ids, const int* row_start, const int* lens, KvGeom kv,
                              int layer, int context, void* out, int out_type, int n, cudaStream_t st, int max_len) {
  if (kv.D != 64) return cudaErrorInvalidValue;
  if (kv.ring && (kv.cap & (kv.cap - 1))) return cudaErrorInvalidValue;
  // chunks longer than one query per warp: an append pass, then an attend pass with one warp per query
  static const bool split_on = [] { const char* e = getenv("PTTS_PREFILL_SPLIT"); return !(e && e[0] == '0'); }();
  const bool split = split_on && max_len > 8 && !kv.ring;
  // bf16 chunks of 16 tokens and more: the attend pass runs on the tensor cores (attn_prefill_mma.cuh); PTTS_PREFILL_MMA=0
  // keeps the one-warp-per-query form
  static const bool mma_on = [] { const char* e = getenv("PTTS_PREFILL_MMA"); return !(e && e[0] == '0'); }();
  const bool mma = mma_on && split && max_len >= 16 && kv.type == 1 && out_type == 1;
  for (int pass = split ? 1 : 0; pass <= (split ? 2 : 0); ++pass) {
    if (pass == 2 && mma) {
      launch_k(attn_prefill_mma_kernel, dim3(kv.H, n, (max_len + 63) / 64), 128, 0, st, (const float*)qkv, offsets, slot_ids, row_start, lens,
               kv, layer, (bf16*)out);
      break;
    }
    dim3 grid(kv.H, n, pass == 2 ? (max_len + 7) / 8 : 1);
    if (kv.type == 0 && out_type == 0)
      launch_k(attn_chunk_kernel<float, float>, grid, 256, 0, st, qkv, rope_freqs, offsets, slot_ids, row_start, lens, kv, layer, context, (float*)out, pass);
    else if (kv.type == 1 && out_type == 1)
      launch_k(attn_chunk_kernel<bf16, bf16>, grid, 256, 0, st, qkv, rope_freqs, offsets, slot_ids, row_start, lens, kv, layer, context, (bf16*)out, pass);
    else if (kv.type == 0 && out_type == 1)
      launch_k(attn_chunk_kernel<float, bf16>, grid, 256, 0, st, qkv, rope_freqs, offsets, slot_ids, row_start, lens, kv, layer, context, (bf16*)out, pass);
    else if (kv.type == 2 && out_type == 1 && !kv.ring && kv.scale)
      launch_k(attn_chunk_kernel<fp8e4, bf16>, grid, 256, 0, st, qkv, rope_freqs, offsets, slot_ids, row_start, lens, kv, layer, context, (bf16*)out, pass);
    else
      return cudaErrorInvalidValue;
  }
  return cudaGetLastError();
}

// =====================================================================================
// window: Mimi decoder attention, T (<= 16) new tokens per stream per frame, sliding window
// `context` over a ring cache (transformer.py:22-29, mimi_transformer.py:140-150).
// grid (H, n), 256 threads.  The <= context+T-1 keys/values a frame can see are staged ONCE in
// shared memory (padded rows: conflict-free both when lanes own keys and when lanes own dims);
// scores: lanes own keys (no shuffles in the dot products), PV: lanes own dim pairs.
// =====================================================================================
template <typename TK> struct KvWords;
template <> struct KvWords<bf16> { static constexpr int ROW = 32, LD = 33; };    // 64 bf16 = 32 words (+1 pad)
template <> struct KvWords<float> { static constexpr int ROW = 64, LD = 65; };

__device__ __forceinline__ float2 kv_word_pair(const uint32_t* row, int c, bf16) {   // dims 2c, 2c+1
  return __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(row + c));
}
__device__ __forceinline__ float2 kv_word_pair(const uint32_t* row, int c, float) {
  return make_float2(__uint_as_float(row[2 * c]), __uint_as_float(row[2 * c + 1]));
}

__device__ __forceinline__ void cp_async4(uint32_t* smem_dst, const uint32_t* gmem_src) {
  asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"((uint32_t)__cvta_generic_to_shared(smem_dst)), "l"(gmem_src) : "memory");
}
template <typename TK> __device__ __forceinline__ void kv_smem_store(uint32_t* row, int lane, float2 v);
template <> __device__ __forceinline__ void kv_smem_store<bf16>(uint32_t* row, int lane, float2 v) {
  const __nv_bfloat162 b = __floats2bfloat162_rn(v.x, v.y);
  row[lane] = *reinterpret_cast<const uint32_t*>(&b);
}
template <> __device__ __forceinline__ void kv_smem_store<float>(uint32_t* row, int lane, float2 v) {
  row[2 * lane] = __float_as_uint(v.x);
  row[2 * lane + 1] = __float_as_uint(v.y);
}

// grid (H, n, frames): block z owns the T tokens of frame z of this pass (a pass appends frames * T tokens
// per stream).  Keys / values of positions before the pass come from the ring; those of this pass (own
// frame and earlier frames) are rotated straight from the packed projection and rounded to the cache
// type, so no block depends on another block's ring writes; a block appends only its own T tokens.
template <typename TK, typename TO>
__global__ void __launch_bounds__(256)
attn_window_kernel(const float* __restrict__ qkv, const float* __restrict__ freqs, const int* __restrict__ offsets,
                   const int* __restrict__ slot_ids, KvGeom g, int layer, int context, int T, int maxw,
                   TO* __restrict__ out) {
  constexpr int NW = 8, TQ = 16, NKL = 9;   // warps, max queries, max keys per lane (maxw <= 288)
  constexpr int LD = KvWords<TK>::LD;
  extern __shared__ __align__(16) uint32_t smw[];
  uint32_t* Ks = smw;
  uint32_t* Vs = Ks + maxw * LD;
  float* Ps = reinterpret_cast<float*>(Vs + maxw * LD);   // [TQ][maxw]
  float* Qs = Ps + TQ * maxw;                              // [TQ][64]
  pdl_launch_dependents();
  pdl_wait();                                              // qkv comes from the previous kernel
  const int h = blockIdx.x, i = blockIdx.y, z = blockIdx.z, frames = gridDim.z;
  const int slot = slot_ids[i];
  const int off0 = offsets[slot];                          // position of the pass's first token
  const int off = off0 + z * T;                            // position of this block's first token
  const int r_pass = i * frames * T;                       // first projection row of the stream's pass
  const int r0 = r_pass + z * T;
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  const int d_model = g.H * 64;
  const float fr = freqs[lane];

  // phase 1: rotated q of the own tokens
  for (int t = w; t < T; t += NW) {
    const float* row = qkv + (long long)(r0 + t) * 3 * d_model + h * 64 + 2 * lane;
    const float2 q = rope_pair(load2(row), (float)(off + t), fr);
    Qs[t * 64 + 2 * lane] = q.x;
    Qs[t * 64 + 2 * lane + 1] = q.y;
  }
  // phase 2: stage the visible window [lo, off+T-1]
  const int lo = max(0, off - context + 1);
  const int W = off + T - lo;
  for (int j = w; j < W; j += NW) {
    const int pj = lo + j;
    if (pj < off0) {
      // asynchronous global -> shared copies: every row of the window is in flight at once (the register
      // path exposed one L2 round trip per few rows)
      const uint32_t* kr = reinterpret_cast<const uint32_t*>(kv_row<TK>(g, layer, slot, 0, h, pj));
      const uint32_t* vr = reinterpret_cast<const uint32_t*>(kv_row<TK>(g, layer, slot, 1, h, pj));
      constexpr int WPL = sizeof(TK) == 2 ? 1 : 2;   // 32-bit words per lane
#pragma unroll
      for (int q = 0; q < WPL; ++q) {
        cp_async4(Ks + j * LD + WPL * lane + q, kr + WPL * lane + q);
        cp_async4(Vs + j * LD + WPL * lane + q, vr + WPL * lane + q);
      }
    } else {
      const float* row = qkv + (long long)(r_pass + (pj - off0)) * 3 * d_model + h * 64 + 2 * lane;
      const float2 k = rope_pair(load2(row + d_model), (float)pj, fr);
      const float2 v = load2(row + 2 * d_model);
      kv_smem_store<TK>(Ks + j * LD, lane, k);
      kv_smem_store<TK>(Vs + j * LD, lane, v);
      if (pj >= off) {   // own token: append to the ring
        store2(kv_row<TK>(g, layer, slot, 0, h, pj) + 2 * lane, k);
        store2(kv_row<TK>(g, layer, slot, 1, h, pj) + 2 * lane, v);
      }
    }
  }
  asm volatile("cp.async.wait_all;" ::: "memory");
  __syncthreads();

  // phase 3: one warp per query
  for (int t = w; t < T; t += NW) {
    const int pq = off + t;
    const int jlo = max(0, pq - context + 1) - lo, jhi = pq - lo;   // inclusive key range in the window
    float q[64];
#pragma unroll
    for (int c = 0; c < 64; c += 4) {
      const float4 t4 = *reinterpret_cast<const float4*>(Qs + t * 64 + c);
      q[c] = t4.x; q[c + 1] = t4.y; q[c + 2] = t4.z; q[c + 3] = t4.w;
    }
    float sc[NKL];
    float mx = -INFINITY;
#pragma unroll
    for (int u = 0; u < NKL; ++u) {
      const int j = jlo + lane + 32 * u;
      float a = -INFINITY;
      if (j <= jhi) {
        const uint32_t* kr = Ks + j * LD;
        float acc = 0.0f;
#pragma unroll
        for (int c = 0; c < 32; ++c) {
          const float2 kk = kv_word_pair(kr, c, TK());
          acc = fmaf(q[2 * c], kk.x, acc);
          acc = fmaf(q[2 * c + 1], kk.y, acc);
        }
        a = acc * 0.125f;
      }
      sc[u] = a;
      mx = fmaxf(mx, a);
    }
    mx = warp_max(mx);
    float l = 0.0f;
#pragma unroll
    for (int u = 0; u < NKL; ++u) {
      const int j = jlo + lane + 32 * u;
      if (j <= jhi) {
        const float p = expf(sc[u] - mx);
        l += p;
        Ps[t * maxw + j] = p;
      }
    }
    l = warp_sum(l);
    __syncwarp();
    float2 acc = make_float2(0.f, 0.f);
    const float* pr = Ps + t * maxw;
    for (int j = jlo; j <= jhi; ++j) {
      const float p = pr[j];
      const float2 vv = kv_word_pair(Vs + j * LD, lane, TK());
      acc.x = fmaf(p, vv.x, acc.x);
      acc.y = fmaf(p, vv.y, acc.y);
    }
    const float inv = 1.0f / l;
    store2(out + (long long)(r0 + t) * d_model + h * 64 + 2 * lane, make_float2(acc.x * inv, acc.y * inv));
  }
}

}  // namespace ptts
namespace ptts {

cudaError_t launch_attn_window(const float* qkv, const float* rope_freqs, const int* offsets, const int* slot_ids,
                               KvGeom kv, int layer, int context, int T, void* out, int out_type, int n,
                               cudaStream_t st, int frames, bool alone) {
  if (kv.D != 64 || !kv.ring || (kv.cap & (kv.cap - 1)) || T < 1 || T > 16 || frames < 1) return cudaErrorInvalidValue;
  if (context + frames * T > kv.cap) return cudaErrorInvalidValue;   // a pass must not wrap onto keys it still reads
  const int maxw = ((context + T - 1 + 7) / 8) * 8;
  if (maxw > 288 || maxw > kv.cap) return cudaErrorInvalidValue;
  const char* env_mma = getenv("PTTS_WIN_MMA");   // read per call: the tests compare both kernels in one process
  const bool use_mma = !env_mma || atoi(env_mma) != 0;
  const int NQ = frames * T;
  if (use_mma && kv.type == 1 && out_type == 1 && T == 16 && NQ <= 64 && context - 1 + NQ <= WM_MAXW) {
    // PTTS_MIMI_HALF=1 (probe): one CTA per SM (shared memory padded past half an SM) so that a kernel of the
    // latency-critical FlowLM branch always finds room beside it
    static const bool half_sm = [] { const char* e = getenv("PTTS_MIMI_HALF"); return e && e[0] == '1'; }();
    // Full passes of large batches stream the window through a double buffer, four CTAs per SM (attn_window_stream_kernel):
    // Mimi pass 0.658 -> 0.634 ms at 64 streams, 7.36 -> 7.21 ms at 1024.  At 64 streams the full frame step does not gain
    // (0.464 vs 0.466 ms: the denser kernel disturbs the FlowLM chain as much as it saves), so small batches keep the staged
    // form.  PTTS_WIN_STREAM=0 / 1 forces one (read per call: the tests compare both).
    const char* env_stream = getenv("PTTS_WIN_STREAM");
    const bool stream_form = env_stream ? env_stream[0] != '0' : (alone || n >= 96);   // alone: no FlowLM chain beside the pass
    if (NQ == 64 && !half_sm && stream_form) {
      static DeviceOnce sattr;
      if (sattr.first()) {
        cudaError_t e = cudaFuncSetAttribute(attn_window_stream_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)attn_window_stream_smem());
        if (e != cudaSuccess) return e;
      }
      launch_k(attn_window_stream_kernel, dim3(kv.H, n), WM_THREADS, attn_window_stream_smem(), st, qkv, rope_freqs, offsets,
               slot_ids, kv, layer, context, (bf16*)out);
      return cudaGetLastError();
    }
    const size_t wsmem = half_sm ? std::max<size_t>(attn_window_mma_smem(), 118 * 1024) : attn_window_mma_smem();
    static DeviceOnce attr;
    if (attr.first()) {
      cudaError_t e = cudaFuncSetAttribute(attn_window_mma_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)wsmem);
      if (e != cudaSuccess) return e;
    }
    launch_k(attn_window_mma_kernel, dim3(kv.H, n), WM_THREADS, wsmem, st, qkv, rope_freqs, offsets,
             slot_ids, kv, layer, context, NQ, (bf16*)out);
    return cudaGetLastError();
  }
  dim3 grid(kv.H, n, frames);
  const int ld = kv.type == 1 ? KvWords<bf16>::LD : KvWords<float>::LD;
  const size_t smem = ((size_t)2 * maxw * ld + (size_t)16 * maxw + 16 * 64) * 4;
#define WIN_CASE(TKT, TOT)                                                                                        \
  do {                                                                                                            \
    static DeviceOnce attr;                                                                                     \
    if (attr.first()) {                                                                                                  \
      cudaError_t e = cudaFuncSetAttribute(attn_window_kernel<TKT, TOT>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024); \
      if (e != cudaSuccess) return e;                                                                             \
                                                                                                                        \
    }                                                                                                             \
    launch_k(attn_window_kernel<TKT, TOT>, grid, 256, smem, st, qkv, rope_freqs, offsets, slot_ids, kv, layer,    \
             context, T, maxw, (TOT*)out);                                                                        \
  } while (0)
  if (kv.type == 0 && out_type == 0) WIN_CASE(float, float);
  else if (kv.type == 1 && out_type == 1) WIN_CASE(bf16, bf16);
  else if (kv.type == 0 && out_type == 1) WIN_CASE(float, bf16);
  else return cudaErrorInvalidValue;
#undef WIN_CASE
  return cudaGetLastError();
}

// =====================================================================================
// KV import / export / fork (model_state dict <-> slot)   tts_model.py:637-638,1047-1072
// reference layout per layer: cache[2, 1, P, H, D]
// =====================================================================================
template <typename TK>
__global__ void kv_import_kernel(const float* __restrict__ ref, KvGeom g, int layer, int slot, int P) {
  pdl_launch_dependents();
  pdl_wait();
  const long long total = 2LL * P * g.H * g.D;
  for (long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x; idx < total;
       idx += (long long)gridDim.x * blockDim.x) {
    const int d = (int)(idx % g.D);
    long long r = idx / g.D;
    const int h = (int)(r % g.H); r /= g.H;
    const int p = (int)(r % P);
    const int kv = (int)(r / P);
    kv_row<TK>(g, layer, slot, kv, h, p)[d] = from_f<TK>(ref[idx]);
  }
}
// positions [p0, p1) of `slot` -> rows [p0, p1) of ref [2, P, H, D]
template <typename TK>
__global__ void kv_export_range_kernel(float* __restrict__ ref, KvGeom g, int layer, int slot, int P, int p0, int p1) {
  pdl_launch_dependents();
  pdl_wait();
  const int np = p1 - p0;
  const long long total = 2LL * np * g.H * g.D;
  for (long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x; idx < total;
       idx += (long long)gridDim.x * blockDim.x) {
    const int d = (int)(idx % g.D);
    long long r = idx / g.D;
    const int h = (int)(r % g.H); r /= g.H;
    const int p = p0 + (int)(r % np);
    const int kv = (int)(r / np);
    ref[(((long long)kv * P + p) * g.H + h) * g.D + d] = to_f(kv_row<TK>(g, layer, slot, kv, h, p)[d]);
  }
}
template <typename TK>
__global__ void kv_export_kernel(float* __restrict__ ref, KvGeom g, int layer, int slot, int P) {
  pdl_launch_dependents();
  pdl_wait();
  const long long total = 2LL * P * g.H * g.D;
  for (long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x; idx < total;
       idx += (long long)gridDim.x * blockDim.x) {
    const int d = (int)(idx % g.D);
    long long r = idx / g.D;
    const int h = (int)(r % g.H); r /= g.H;
    const int p = (int)(r % P);
    const int kv = (int)(r / P);
    ref[idx] = to_f(kv_row<TK>(g, layer, slot, kv, h, p)[d]);
  }
}
template <typename TK>
__global__ void kv_fork_kernel(KvGeom g, int n_layers, int src, int dst, int p0, int P) {
  pdl_launch_dependents();
  pdl_wait();
  // copy positions [p0,P) of every (layer, kv, head): (P - p0)*D contiguous elements each
  const int seg = blockIdx.x;  // layer*2*H + kv*H + h
  const int h = seg % g.H, kv = (seg / g.H) % 2, layer = seg / (2 * g.H);
  const TK* s = kv_row<TK>(g, layer, src, kv, h, p0);
  TK* d = kv_row<TK>(g, layer, dst, kv, h, p0);
  const int total = (P - p0) * g.D;
  for (int idx = threadIdx.x * 4; idx < total; idx += blockDim.x * 4) store4(d + idx, load4(s + idx));
}

// ---- FP8 cache: one warp per (k|v, position, head) row
__global__ void kv_import_fp8_kernel(const float* __restrict__ ref, KvGeom g, int layer, int slot, int P) {
  pdl_launch_dependents();
  pdl_wait();
  const int lane = threadIdx.x & 31;
  const long long rows = 2LL * P * g.H;
  for (long long r = (long long)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5); r < rows; r += (long long)gridDim.x * (blockDim.x >> 5)) {
    const int h = (int)(r % g.H), p = (int)((r / g.H) % P), kv = (int)(r / ((long long)g.H * P));
    float2 v = load2(ref + r * 64 + 2 * lane);
    const float sc = fp8_scale(warp_max(fmaxf(fabsf(v.x), fabsf(v.y))));
    v.x /= sc; v.y /= sc;
    store2(kv_row<fp8e4>(g, layer, slot, kv, h, p) + 2 * lane, v);
    if (lane == 0) *kv_scale_row(g, layer, slot, kv, h, p) = sc;
  }
}
__global__ void kv_export_fp8_kernel(float* __restrict__ ref, KvGeom g, int layer, int slot, int P, int p0, int p1) {
  pdl_launch_dependents();
  pdl_wait();
  const int lane = threadIdx.x & 31, np = p1 - p0;
  const long long rows = 2LL * np * g.H;
  for (long long r = (long long)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5); r < rows; r += (long long)gridDim.x * (blockDim.x >> 5)) {
    const int h = (int)(r % g.H), p = p0 + (int)((r / g.H) % np), kv = (int)(r / ((long long)g.H * np));
    float2 v = load2(kv_row<fp8e4>(g, layer, slot, kv, h, p) + 2 * lane);
    const float sc = *kv_scale_row(g, layer, slot, kv, h, p);
    store2(ref + (((long long)kv * P + p) * g.H + h) * 64 + 2 * lane, make_float2(v.x * sc, v.y * sc));
  }
}
__global__ void kv_fork_fp8_kernel(KvGeom g, int src, int dst, int p0, int p1) {
  pdl_launch_dependents();
  pdl_wait();
  const int seg = blockIdx.x;  // layer*2*H + kv*H + h
  const int h = seg % g.H, kv = (seg / g.H) % 2, layer = seg / (2 * g.H);
  const uint32_t* s = reinterpret_cast<const uint32_t*>(kv_row<fp8e4>(g, layer, src, kv, h, p0));
  uint32_t* d = reinterpret_cast<uint32_t*>(kv_row<fp8e4>(g, layer, dst, kv, h, p0));
  for (int idx = threadIdx.x; idx < (p1 - p0) * 16; idx += blockDim.x) d[idx] = s[idx];
  const float* ss = kv_scale_row(g, layer, src, kv, h, p0);
  float* ds = kv_scale_row(g, layer, dst, kv, h, p0);
  for (int idx = threadIdx.x; idx < p1 - p0; idx += blockDim.x) ds[idx] = ss[idx];
}

cudaError_t launch_kv_import(const float* ref, KvGeom kv, int layer, int slot, int P, cudaStream_t st) {
  if (P <= 0) return cudaSuccess;
  if (kv.type == 2) {
    launch_k(kv_import_fp8_kernel, (int)min(1024LL, (2LL * P * kv.H + 7) / 8), 256, 0, st, ref, kv, layer, slot, P);
    return cudaGetLastError();
  }
  const int blocks = (int)min(1024LL, (2LL * P * kv.H * kv.D + 255) / 256);
  if (kv.type == 0) launch_k(kv_import_kernel<float>, blocks, 256, 0, st, ref, kv, layer, slot, P);
  else launch_k(kv_import_kernel<bf16>, blocks, 256, 0, st, ref, kv, layer, slot, P);
  return cudaGetLastError();
}
cudaError_t launch_kv_export(float* ref, KvGeom kv, int layer, int slot, int P, cudaStream_t st) {
  if (P <= 0) return cudaSuccess;
  if (kv.type == 2) return launch_kv_export_range(ref, kv, layer, slot, P, 0, P, st);
  const int blocks = (int)min(1024LL, (2LL * P * kv.H * kv.D + 255) / 256);
  if (kv.type == 0) launch_k(kv_export_kernel<float>, blocks, 256, 0, st, ref, kv, layer, slot, P);
  else launch_k(kv_export_kernel<bf16>, blocks, 256, 0, st, ref, kv, layer, slot, P);
  return cudaGetLastError();
}
cudaError_t launch_kv_export_range(float* ref, KvGeom kv, int layer, int slot, int P, int p0, int p1, cudaStream_t st) {
  if (p1 <= p0) return cudaSuccess;
  if (kv.type == 2) {
    launch_k(kv_export_fp8_kernel, (int)min(1024LL, (2LL * (p1 - p0) * kv.H + 7) / 8), 256, 0, st, ref, kv, layer, slot, P, p0, p1);
    return cudaGetLastError();
  }
  const int blocks = (int)min(1024LL, (2LL * (p1 - p0) * kv.H * kv.D + 255) / 256);
  if (kv.type == 0) launch_k(kv_export_range_kernel<float>, blocks, 256, 0, st, ref, kv, layer, slot, P, p0, p1);
  else launch_k(kv_export_range_kernel<bf16>, blocks, 256, 0, st, ref, kv, layer, slot, P, p0, p1);
  return cudaGetLastError();
}
// positions [p0, p1) of every (layer, k|v, head) of src -> dst
cudaError_t launch_kv_fork_range(KvGeom kv, int n_layers, int src, int dst, int p0, int p1, cudaStream_t st);
cudaError_t launch_kv_fork(KvGeom kv, int n_layers, int src, int dst, int P, cudaStream_t st) {
  if (P <= 0) return cudaSuccess;
  if (kv.type == 2) return launch_kv_fork_range(kv, n_layers, src, dst, 0, P, st);
  const int blocks = n_layers * 2 * kv.H;
  if (kv.type == 0) launch_k(kv_fork_kernel<float>, blocks, 256, 0, st, kv, n_layers, src, dst, 0, P);
  else launch_k(kv_fork_kernel<bf16>, blocks, 256, 0, st, kv, n_layers, src, dst, 0, P);
  return cudaGetLastError();
}
cudaError_t launch_kv_fork_range(KvGeom kv, int n_layers, int src, int dst, int p0, int p1, cudaStream_t st) {
  if (p1 <= p0) return cudaSuccess;
  const int blocks = n_layers * 2 * kv.H;
  if (kv.type == 2) {
    launch_k(kv_fork_fp8_kernel, blocks, 256, 0, st, kv, src, dst, p0, p1);
    return cudaGetLastError();
  }
  if (kv.type == 0) launch_k(kv_fork_kernel<float>, blocks, 256, 0, st, kv, n_layers, src, dst, p0, p1);
  else launch_k(kv_fork_kernel<bf16>, blocks, 256, 0, st, kv, n_layers, src, dst, p0, p1);
  return cudaGetLastError();
}

}  // namespace ptts
