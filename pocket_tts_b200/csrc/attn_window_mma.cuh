// Mimi decoder window attention on the warp-level tensor-core path (bf16 cache, bf16 output).
// Included by attention.cu (uses kv_row / rope_pair / load2 from there).
//
// One CTA per (stream, head) owns every query of a pass (NQ = frames * T <= 64 tokens, a multiple of 16):
// the <= context - 1 + NQ visible keys / values are staged once in shared memory (16 B asynchronous copies
// from the ring, the pass's own tokens rotated straight from the packed projection), then the four warps
// run a flash-style loop over 64-key blocks with mma.sync.m16n8k16 (S = Q K^T and O = P V, fp32
// accumulators, P rounded to bf16).  Warp w owns query tile (w % qtiles) and key split (w / qtiles), so a
// single-frame pass (16 queries) still uses all four warps; splits are merged by log-sum-exp through
// shared memory.  transformer.py:22-29,135-158, mimi_transformer.py:140-150.
#pragma once

namespace ptts {

constexpr int WM_LD = 72;        // bf16 per staged row (64 + 8 pad: ldmatrix rows fall in distinct 16 B banks)
constexpr int WM_MAXW = 320;     // staged rows (context 250 - 1 + 64 = 313, padded to the 64-key block)
constexpr int WM_THREADS = 128;

__device__ __forceinline__ void ldsm_x4(uint32_t (&r)[4], const bf16* p) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0,%1,%2,%3}, [%4];"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]) : "r"((uint32_t)__cvta_generic_to_shared(p)));
}
__device__ __forceinline__ void ldsm_x4_t(uint32_t (&r)[4], const bf16* p) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.trans.shared.b16 {%0,%1,%2,%3}, [%4];"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]) : "r"((uint32_t)__cvta_generic_to_shared(p)));
}
__device__ __forceinline__ void mma_bf16_16816(float (&c)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
  asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
               : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
               : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}
__device__ __forceinline__ uint32_t pack_bf16(float lo, float hi) {
  const __nv_bfloat162 b = __floats2bfloat162_rn(lo, hi);
  return *reinterpret_cast<const uint32_t*>(&b);
}
__device__ __forceinline__ void cp_async16(void* smem_dst, const void* gmem_src) {
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"((uint32_t)__cvta_generic_to_shared(smem_dst)), "l"(gmem_src) : "memory");
}

__global__ void __launch_bounds__(WM_THREADS, 2)
attn_window_mma_kernel(const float* __restrict__ qkv, const float* __restrict__ freqs, const int* __restrict__ offsets,
                       const int* __restrict__ slot_ids, KvGeom g, int layer, int context, int NQ,
                       bf16* __restrict__ out) {
  extern __shared__ __align__(16) unsigned char wm_smem[];
  bf16* Qs = reinterpret_cast<bf16*>(wm_smem);   // [64][WM_LD]
  bf16* Ks = Qs + 64 * WM_LD;                    // [WM_MAXW][WM_LD]
  bf16* Vs = Ks + WM_MAXW * WM_LD;
  pdl_launch_dependents();
  pdl_wait();                                    // qkv comes from the previous kernel
  const int h = blockIdx.x, i = blockIdx.y;
  const int slot = slot_ids[i];
  const int off0 = offsets[slot];                // position of the pass's first token
  const int r_pass = i * NQ;                     // first projection row of the stream's pass
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  const int d_model = g.H * 64;
  const int lo = max(0, off0 - context + 1);
  const int n_ring = off0 - lo;                  // staged rows that come from the ring
  const int W = n_ring + NQ;                     // staged rows in all
  const int Wp = (W + 63) & ~63;

  // ---- stage: ring rows by 16 B asynchronous copies (8 per row, K and V)
  for (int idx = threadIdx.x; idx < n_ring * 8; idx += WM_THREADS) {
    const int j = idx >> 3, c = idx & 7;
    cp_async16(Ks + j * WM_LD + c * 8, kv_row<bf16>(g, layer, slot, 0, h, lo + j) + c * 8);
    cp_async16(Vs + j * WM_LD + c * 8, kv_row<bf16>(g, layer, slot, 1, h, lo + j) + c * 8);
  }
  // ---- the pass's own tokens: rotate, round, stage, append to the ring; q rotated and staged
  {
    const float fr = freqs[lane];
    // a warp takes tokens w, w + 4, ...: the three projection rows of FOUR tokens are requested before the first is
    // used (one token per iteration was 16 serial global round trips per CTA at 64 tokens, the longest phase of the kernel)
    constexpr int TB = 4, WARPS = WM_THREADS / 32;
    for (int t0 = w; t0 < NQ; t0 += TB * WARPS) {
      float2 qv[TB], kv[TB], vv[TB];
#pragma unroll
      for (int u = 0; u < TB; ++u) {
        const int t = t0 + u * WARPS;
        if (t < NQ) {
          const float* row = qkv + (long long)(r_pass + t) * 3 * d_model + h * 64 + 2 * lane;
          qv[u] = load2(row); kv[u] = load2(row + d_model); vv[u] = load2(row + 2 * d_model);
        }
      }
#pragma unroll
      for (int u = 0; u < TB; ++u) {
        const int t = t0 + u * WARPS;
        if (t < NQ) {
          const float pos = (float)(off0 + t);
          const float2 q = rope_pair(qv[u], pos, fr);
          const float2 k = rope_pair(kv[u], pos, fr);
          const uint32_t kb = pack_bf16(k.x, k.y), vb = pack_bf16(vv[u].x, vv[u].y);
          reinterpret_cast<uint32_t*>(Qs + t * WM_LD)[lane] = pack_bf16(q.x, q.y);
          reinterpret_cast<uint32_t*>(Ks + (n_ring + t) * WM_LD)[lane] = kb;
          reinterpret_cast<uint32_t*>(Vs + (n_ring + t) * WM_LD)[lane] = vb;
          reinterpret_cast<uint32_t*>(kv_row<bf16>(g, layer, slot, 0, h, off0 + t))[lane] = kb;
          reinterpret_cast<uint32_t*>(kv_row<bf16>(g, layer, slot, 1, h, off0 + t))[lane] = vb;
        }
      }
    }
    // rows of the last key block past the window: finite values (their probabilities are zero)
    for (int idx = threadIdx.x; idx < (Wp - W) * 32; idx += WM_THREADS) {
      const int j = W + (idx >> 5), c = idx & 31;
      reinterpret_cast<uint32_t*>(Ks + j * WM_LD)[c] = 0u;
      reinterpret_cast<uint32_t*>(Vs + j * WM_LD)[c] = 0u;
    }
  }
  asm volatile("cp.async.wait_all;" ::: "memory");
  __syncthreads();

  // ---- warp roles
  const int qtiles = NQ >> 4, splits = (WM_THREADS / 32) / qtiles;   // (4,1) (2,2) (1,4); NQ = 48: (3,1) + idle warp
  const int qt = w % qtiles, sp = w / qtiles;
  const bool active = sp < splits;
  const int gq = lane >> 2, qq = lane & 3;
  float o[8][4];
#pragma unroll
  for (int nt = 0; nt < 8; ++nt) o[nt][0] = o[nt][1] = o[nt][2] = o[nt][3] = 0.0f;
  float m_run[2] = {-INFINITY, -INFINITY}, l_run[2] = {0.0f, 0.0f};
  if (active) {
    // visible staged-row range of this query tile, in 64-key blocks, divided among the splits
    const int p_first = off0 + 16 * qt, p_last = p_first + 15;
    const int jmin = max(0, p_first - context + 1 - lo), jmax = p_last - lo;
    const int b_lo = jmin >> 6, b_hi = (jmax >> 6) + 1;
    const int nb = b_hi - b_lo, per = (nb + splits - 1) / splits;
    const int kb_beg = b_lo + sp * per, kb_end = min(b_hi, kb_beg + per);
    uint32_t qa[4][4];
#pragma unroll
    for (int kk = 0; kk < 4; ++kk) ldsm_x4(qa[kk], Qs + (16 * qt + (lane & 15)) * WM_LD + kk * 16 + (lane >> 4) * 8);
    const int pq0 = p_first + gq, pq1 = pq0 + 8;   // the two query positions this thread holds
    for (int kb = kb_beg; kb < kb_end; ++kb) {
      float s[8][4];
#pragma unroll
      for (int nt = 0; nt < 8; ++nt) s[nt][0] = s[nt][1] = s[nt][2] = s[nt][3] = 0.0f;
#pragma unroll
      for (int kk = 0; kk < 4; ++kk) {
#pragma unroll
        for (int np = 0; np < 4; ++np) {   // pairs of 8-key tiles
          uint32_t b[4];
          const int key = kb * 64 + np * 16 + ((lane >> 4) << 3) + (lane & 7);
          ldsm_x4(b, Ks + key * WM_LD + kk * 16 + ((lane >> 3) & 1) * 8);
          mma_bf16_16816(s[2 * np], qa[kk], b[0], b[1]);
          mma_bf16_16816(s[2 * np + 1], qa[kk], b[2], b[3]);
        }
      }
      // scale, mask (pq - context < pj <= pq), block row maxima
      float mx[2] = {-INFINITY, -INFINITY};
#pragma unroll
      for (int nt = 0; nt < 8; ++nt) {
#pragma unroll
        for (int e = 0; e < 4; ++e) {
          const int pj = lo + kb * 64 + nt * 8 + 2 * qq + (e & 1);
          const int pq = (e < 2) ? pq0 : pq1;
          const bool vis = pj <= pq && pj > pq - context;
          const float v = vis ? s[nt][e] * 0.125f : -INFINITY;
          s[nt][e] = v;
          mx[e >> 1] = fmaxf(mx[e >> 1], v);
        }
      }
      float corr[2], m_use[2];
#pragma unroll
      for (int r = 0; r < 2; ++r) {
        mx[r] = fmaxf(mx[r], __shfl_xor_sync(0xffffffffu, mx[r], 1));
        mx[r] = fmaxf(mx[r], __shfl_xor_sync(0xffffffffu, mx[r], 2));
        const float m_new = fmaxf(m_run[r], mx[r]);
        m_use[r] = m_new == -INFINITY ? 0.0f : m_new;
        corr[r] = __expf(m_run[r] - m_use[r]);     // first block: exp(-inf) = 0
        m_run[r] = m_new;
        l_run[r] *= corr[r];
      }
#pragma unroll
      for (int nt = 0; nt < 8; ++nt) {
        o[nt][0] *= corr[0]; o[nt][1] *= corr[0];
        o[nt][2] *= corr[1]; o[nt][3] *= corr[1];
      }
      uint32_t pa[4][4];
#pragma unroll
      for (int nt = 0; nt < 8; ++nt) {
        const float p0 = __expf(s[nt][0] - m_use[0]), p1 = __expf(s[nt][1] - m_use[0]);
        const float p2 = __expf(s[nt][2] - m_use[1]), p3 = __expf(s[nt][3] - m_use[1]);
        l_run[0] += p0 + p1;
        l_run[1] += p2 + p3;
        pa[nt >> 1][(nt & 1) * 2] = pack_bf16(p0, p1);
        pa[nt >> 1][(nt & 1) * 2 + 1] = pack_bf16(p2, p3);
      }
#pragma unroll
      for (int kk = 0; kk < 4; ++kk) {     // 16 keys
#pragma unroll
        for (int np = 0; np < 4; ++np) {   // pairs of 8-dim tiles
          uint32_t b[4];
          const int key = kb * 64 + kk * 16 + (((lane >> 3) & 1) << 3) + (lane & 7);
          ldsm_x4_t(b, Vs + key * WM_LD + np * 16 + (lane >> 4) * 8);
          mma_bf16_16816(o[2 * np], pa[kk], b[0], b[1]);
          mma_bf16_16816(o[2 * np + 1], pa[kk], b[2], b[3]);
        }
      }
    }
#pragma unroll
    for (int r = 0; r < 2; ++r) {          // row sums live in the four lanes of a quad
      l_run[r] += __shfl_xor_sync(0xffffffffu, l_run[r], 1);
      l_run[r] += __shfl_xor_sync(0xffffffffu, l_run[r], 2);
    }
  }

  if (splits > 1) {
    // ---- merge the key splits of a query tile: (m, l, O) of every warp through shared memory
    __syncthreads();                       // every warp is done with the staged keys
    float* Ms = reinterpret_cast<float*>(Ks);               // [4 warps][16 rows][2]  (m, l)
    float* Os = Ms + 4 * 32;                                // [4 warps][16][64]
    if (active) {
      if (qq == 0) {
        Ms[w * 32 + gq * 2] = m_run[0];       Ms[w * 32 + gq * 2 + 1] = l_run[0];
        Ms[w * 32 + (gq + 8) * 2] = m_run[1]; Ms[w * 32 + (gq + 8) * 2 + 1] = l_run[1];
      }
#pragma unroll
      for (int nt = 0; nt < 8; ++nt) {
        *reinterpret_cast<float2*>(Os + w * 1024 + gq * 64 + nt * 8 + 2 * qq) = make_float2(o[nt][0], o[nt][1]);
        *reinterpret_cast<float2*>(Os + w * 1024 + (gq + 8) * 64 + nt * 8 + 2 * qq) = make_float2(o[nt][2], o[nt][3]);
      }
    }
    __syncthreads();
    // all 128 threads: (query row, 8-dim group) items of every query tile
    for (int item = threadIdx.x; item < NQ * 8; item += WM_THREADS) {
      const int row = item >> 3, dg = item & 7, tq = row >> 4, r = row & 15;
      float m = -INFINITY;
      for (int s2 = 0; s2 < splits; ++s2) m = fmaxf(m, Ms[(s2 * qtiles + tq) * 32 + r * 2]);
      float l = 0.0f, acc[8];
#pragma unroll
      for (int c = 0; c < 8; ++c) acc[c] = 0.0f;
      for (int s2 = 0; s2 < splits; ++s2) {
        const int ww = s2 * qtiles + tq;
        const float ms = Ms[ww * 32 + r * 2];
        const float f = ms == -INFINITY ? 0.0f : __expf(ms - m);
        l += f * Ms[ww * 32 + r * 2 + 1];
        const float4 a = *reinterpret_cast<const float4*>(Os + ww * 1024 + r * 64 + dg * 8);
        const float4 b = *reinterpret_cast<const float4*>(Os + ww * 1024 + r * 64 + dg * 8 + 4);
        acc[0] += f * a.x; acc[1] += f * a.y; acc[2] += f * a.z; acc[3] += f * a.w;
        acc[4] += f * b.x; acc[5] += f * b.y; acc[6] += f * b.z; acc[7] += f * b.w;
      }
      const float inv = 1.0f / l;
      uint4 pk;
      pk.x = pack_bf16(acc[0] * inv, acc[1] * inv); pk.y = pack_bf16(acc[2] * inv, acc[3] * inv);
      pk.z = pack_bf16(acc[4] * inv, acc[5] * inv); pk.w = pack_bf16(acc[6] * inv, acc[7] * inv);
      *reinterpret_cast<uint4*>(out + (long long)(r_pass + row) * d_model + h * 64 + dg * 8) = pk;
    }
  } else if (active) {
    const float inv0 = 1.0f / l_run[0], inv1 = 1.0f / l_run[1];
    bf16* o0 = out + (long long)(r_pass + 16 * qt + gq) * d_model + h * 64 + 2 * qq;
    bf16* o1 = o0 + 8LL * d_model;
#pragma unroll
    for (int nt = 0; nt < 8; ++nt) {
      *reinterpret_cast<uint32_t*>(o0 + nt * 8) = pack_bf16(o[nt][0] * inv0, o[nt][1] * inv0);
      *reinterpret_cast<uint32_t*>(o1 + nt * 8) = pack_bf16(o[nt][2] * inv1, o[nt][3] * inv1);
    }
  }
}

// ---------------------------------------------------------------------------------------------------------------------
// Streaming form for full passes (NQ == 64: four 16-query tiles, one per warp, no key splits).  The staged form above
// keeps the whole window in shared memory (101 KB: two CTAs per SM, 512 CTAs = 1.7 waves at 64 streams).  Here the
// pass's own keys / values go to the ring first and the CTA then walks the visible range lo .. off0 + 63 in 64-key
// blocks through a double buffer (16-byte asynchronous copies of block b + 1 beside the mma work on block b):
// 46 KB and <= 128 registers, four CTAs per SM, the whole grid resident at once.
constexpr int WS_STAGE = 64 * WM_LD;   // bf16 elements of one K (or V) block

__global__ void __launch_bounds__(WM_THREADS, 4)
attn_window_stream_kernel(const float* __restrict__ qkv, const float* __restrict__ freqs, const int* __restrict__ offsets,
                          const int* __restrict__ slot_ids, KvGeom g, int layer, int context, bf16* __restrict__ out) {
  constexpr int NQ = 64;
  extern __shared__ __align__(16) unsigned char wm_smem[];
  bf16* Qs = reinterpret_cast<bf16*>(wm_smem);   // [64][WM_LD]
  bf16* KV = Qs + 64 * WM_LD;                    // [2 buffers][K | V][64][WM_LD]
  pdl_launch_dependents();
  pdl_wait();                                    // qkv comes from the previous kernel
  const int h = blockIdx.x, i = blockIdx.y;
  const int slot = slot_ids[i];
  const int off0 = offsets[slot];
  const int r_pass = i * NQ;
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  const int d_model = g.H * 64;
  const int lo = max(0, off0 - context + 1);
  const int W = off0 - lo + NQ;                  // visible rows: positions lo .. off0 + 63
  const int nblk = (W + 63) >> 6;

  // ---- the pass's own tokens: rotate, round, append to the ring; q rotated and staged
  {
    const float fr = freqs[lane];
    constexpr int TB = 4, WARPS = WM_THREADS / 32;
    for (int t0 = w; t0 < NQ; t0 += TB * WARPS) {
      float2 qv[TB], kv[TB], vv[TB];
#pragma unroll
      for (int u = 0; u < TB; ++u) {
        const float* row = qkv + (long long)(r_pass + t0 + u * WARPS) * 3 * d_model + h * 64 + 2 * lane;
        qv[u] = load2(row); kv[u] = load2(row + d_model); vv[u] = load2(row + 2 * d_model);
      }
#pragma unroll
      for (int u = 0; u < TB; ++u) {
        const int t = t0 + u * WARPS;
        const float pos = (float)(off0 + t);
        const float2 q = rope_pair(qv[u], pos, fr);
        const float2 k = rope_pair(kv[u], pos, fr);
        reinterpret_cast<uint32_t*>(Qs + t * WM_LD)[lane] = pack_bf16(q.x, q.y);
        reinterpret_cast<uint32_t*>(kv_row<bf16>(g, layer, slot, 0, h, off0 + t))[lane] = pack_bf16(k.x, k.y);
        reinterpret_cast<uint32_t*>(kv_row<bf16>(g, layer, slot, 1, h, off0 + t))[lane] = pack_bf16(vv[u].x, vv[u].y);
      }
    }
  }
  __syncthreads();   // the appended rows are read back below by this CTA's own asynchronous copies (L2)

  auto load_block = [&](int b) {
    bf16* Kb = KV + (b & 1) * 2 * WS_STAGE;
    bf16* Vb = Kb + WS_STAGE;
    for (int idx = threadIdx.x; idx < 64 * 8; idx += WM_THREADS) {
      const int j = idx >> 3, c = idx & 7, row = b * 64 + j;
      if (row < W) {
        cp_async16(Kb + j * WM_LD + c * 8, kv_row<bf16>(g, layer, slot, 0, h, lo + row) + c * 8);
        cp_async16(Vb + j * WM_LD + c * 8, kv_row<bf16>(g, layer, slot, 1, h, lo + row) + c * 8);
      } else {   // past the window: finite values (their probabilities are zero)
        *reinterpret_cast<uint4*>(Kb + j * WM_LD + c * 8) = make_uint4(0u, 0u, 0u, 0u);
        *reinterpret_cast<uint4*>(Vb + j * WM_LD + c * 8) = make_uint4(0u, 0u, 0u, 0u);
      }
    }
    asm volatile("cp.async.commit_group;" ::: "memory");
  };

  const int qt = w;
  const int gq = lane >> 2, qq = lane & 3;
  float o[8][4];
#pragma unroll
  for (int nt = 0; nt < 8; ++nt) o[nt][0] = o[nt][1] = o[nt][2] = o[nt][3] = 0.0f;
  float m_run[2] = {-INFINITY, -INFINITY}, l_run[2] = {0.0f, 0.0f};
  const int p_first = off0 + 16 * qt, p_last = p_first + 15;
  const int jmin = max(0, p_first - context + 1 - lo), jmax = p_last - lo;
  const int b_lo = jmin >> 6, b_hi = (jmax >> 6) + 1;       // this warp's visible blocks
  uint32_t qa[4][4];
#pragma unroll
  for (int kk = 0; kk < 4; ++kk) ldsm_x4(qa[kk], Qs + (16 * qt + (lane & 15)) * WM_LD + kk * 16 + (lane >> 4) * 8);
  const int pq0 = p_first + gq, pq1 = pq0 + 8;

  load_block(0);
  for (int kb = 0; kb < nblk; ++kb) {
    if (kb + 1 < nblk) {
      load_block(kb + 1);
      asm volatile("cp.async.wait_group 1;" ::: "memory");
    } else {
      asm volatile("cp.async.wait_group 0;" ::: "memory");
    }
    __syncthreads();
    if (kb >= b_lo && kb < b_hi) {
      const bf16* Ks = KV + (kb & 1) * 2 * WS_STAGE;
      const bf16* Vs = Ks + WS_STAGE;
      float s[8][4];
#pragma unroll
      for (int nt = 0; nt < 8; ++nt) s[nt][0] = s[nt][1] = s[nt][2] = s[nt][3] = 0.0f;
#pragma unroll
      for (int kk = 0; kk < 4; ++kk) {
#pragma unroll
        for (int np = 0; np < 4; ++np) {
          uint32_t b[4];
          const int key = np * 16 + ((lane >> 4) << 3) + (lane & 7);
          ldsm_x4(b, Ks + key * WM_LD + kk * 16 + ((lane >> 3) & 1) * 8);
          mma_bf16_16816(s[2 * np], qa[kk], b[0], b[1]);
          mma_bf16_16816(s[2 * np + 1], qa[kk], b[2], b[3]);
        }
      }
      float mx[2] = {-INFINITY, -INFINITY};
#pragma unroll
      for (int nt = 0; nt < 8; ++nt) {
#pragma unroll
        for (int e = 0; e < 4; ++e) {
          const int pj = lo + kb * 64 + nt * 8 + 2 * qq + (e & 1);
          const int pq = (e < 2) ? pq0 : pq1;
          const bool vis = pj <= pq && pj > pq - context;
          const float v = vis ? s[nt][e] * 0.125f : -INFINITY;
          s[nt][e] = v;
          mx[e >> 1] = fmaxf(mx[e >> 1], v);
        }
      }
      float corr[2], m_use[2];
#pragma unroll
      for (int r = 0; r < 2; ++r) {
        mx[r] = fmaxf(mx[r], __shfl_xor_sync(0xffffffffu, mx[r], 1));
        mx[r] = fmaxf(mx[r], __shfl_xor_sync(0xffffffffu, mx[r], 2));
        const float m_new = fmaxf(m_run[r], mx[r]);
        m_use[r] = m_new == -INFINITY ? 0.0f : m_new;
        corr[r] = __expf(m_run[r] - m_use[r]);
        m_run[r] = m_new;
        l_run[r] *= corr[r];
      }
#pragma unroll
      for (int nt = 0; nt < 8; ++nt) {
        o[nt][0] *= corr[0]; o[nt][1] *= corr[0];
        o[nt][2] *= corr[1]; o[nt][3] *= corr[1];
      }
      uint32_t pa[4][4];
#pragma unroll
      for (int nt = 0; nt < 8; ++nt) {
        const float p0 = __expf(s[nt][0] - m_use[0]), p1 = __expf(s[nt][1] - m_use[0]);
        const float p2 = __expf(s[nt][2] - m_use[1]), p3 = __expf(s[nt][3] - m_use[1]);
        l_run[0] += p0 + p1;
        l_run[1] += p2 + p3;
        pa[nt >> 1][(nt & 1) * 2] = pack_bf16(p0, p1);
        pa[nt >> 1][(nt & 1) * 2 + 1] = pack_bf16(p2, p3);
      }
#pragma unroll
      for (int kk = 0; kk < 4; ++kk) {
#pragma unroll
        for (int np = 0; np < 4; ++np) {
          uint32_t b[4];
          const int key = kk * 16 + (((lane >> 3) & 1) << 3) + (lane & 7);
          ldsm_x4_t(b, Vs + key * WM_LD + np * 16 + (lane >> 4) * 8);
          mma_bf16_16816(o[2 * np], pa[kk], b[0], b[1]);
          mma_bf16_16816(o[2 * np + 1], pa[kk], b[2], b[3]);
        }
      }
    }
    __syncthreads();   // every warp is done with this buffer before block kb + 2 lands in it
  }
#pragma unroll
  for (int r = 0; r < 2; ++r) {
    l_run[r] += __shfl_xor_sync(0xffffffffu, l_run[r], 1);
    l_run[r] += __shfl_xor_sync(0xffffffffu, l_run[r], 2);
  }
  const float inv0 = 1.0f / l_run[0], inv1 = 1.0f / l_run[1];
  bf16* o0 = out + (long long)(r_pass + 16 * qt + gq) * d_model + h * 64 + 2 * qq;
  bf16* o1 = o0 + 8LL * d_model;
#pragma unroll
  for (int nt = 0; nt < 8; ++nt) {
    *reinterpret_cast<uint32_t*>(o0 + nt * 8) = pack_bf16(o[nt][0] * inv0, o[nt][1] * inv0);
    *reinterpret_cast<uint32_t*>(o1 + nt * 8) = pack_bf16(o[nt][2] * inv1, o[nt][3] * inv1);
  }
}

inline size_t attn_window_stream_smem() { return (size_t)(64 + 4 * 64) * WM_LD * sizeof(bf16); }

inline size_t attn_window_mma_smem() { return (size_t)(64 + 2 * WM_MAXW) * WM_LD * sizeof(bf16); }

}  // namespace ptts
