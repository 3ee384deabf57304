// FlowLM prefill attention on the warp-level tensor-core path (bf16 cache, bf16 output): the attend pass of a chunk
// (text tokens or the voice prompt) after the rotate / append pass of attn_chunk_kernel has put the chunk's rotated
// K / V rows into the cache and rotated q in place.  Included by attention.cu after attn_window_mma.cuh (shares its
// ldmatrix / mma.sync helpers).  transformer.py:135-158 with T > 1 (tts_model.py:722-725, 896-899).
//
// One CTA per (head, stream, tile of 64 queries); warp w owns queries 16 w .. 16 w + 15 of the tile.  Keys / values
// of positions [0, off + last query of the tile] stream through a double buffer of 64-key blocks (16-byte cp.async,
// rows of a shared voice prefix come from their owner's cache), flash-style online softmax with mma.sync.m16n8k16 for
// S = Q K^T and O = P V (fp32 accumulators, P rounded to bf16), causal mask pj <= pq.  46 KB of shared memory.
#pragma once

namespace ptts {

constexpr int PM_LD = 72;   // bf16 per staged row (64 + 8 pad)

__global__ void __launch_bounds__(128)
attn_prefill_mma_kernel(const float* __restrict__ qkv, const int* __restrict__ offsets, const int* __restrict__ slot_ids,
                        const int* __restrict__ row_start, const int* __restrict__ lens, KvGeom g, int layer,
                        bf16* __restrict__ out) {
  __shared__ __align__(16) bf16 Qs[64 * PM_LD];
  __shared__ __align__(16) bf16 Kb[2][64 * PM_LD];
  __shared__ __align__(16) bf16 Vb[2][64 * PM_LD];
  pdl_launch_dependents();
  pdl_wait();
  const int h = blockIdx.x, i = blockIdx.y, z = blockIdx.z;
  const int T = lens[i];
  const int t0 = z * 64;
  if (t0 >= T) return;
  const int slot = slot_ids[i], off = offsets[slot], r0 = row_start[i];
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  const int d_model = g.H * 64;
  int plen = 0;
  const long long dsh = pfx_delta(g, slot, plen);
  const int nq = min(64, T - t0);                       // queries of this tile
  const int n_keys = off + t0 + nq;                     // positions [0, n_keys) are visible to the tile's last query
  const int n_kb = (n_keys + 63) >> 6;
  const bf16* kbase = kv_row<bf16>(g, layer, slot, 0, h, 0);
  const bf16* vbase = kv_row<bf16>(g, layer, slot, 1, h, 0);

  auto stage = [&](int kb, int buf) {                   // 64 rows x 8 chunks of K and of V
    for (int idx = threadIdx.x; idx < 64 * 8; idx += 128) {
      const int j = idx >> 3, c = idx & 7, pos = kb * 64 + j;
      bf16* kd = &Kb[buf][j * PM_LD + c * 8];
      bf16* vd = &Vb[buf][j * PM_LD + c * 8];
      if (pos < n_keys) {
        const long long o = (long long)pos * 64 + c * 8 + (pos < plen ? dsh : 0);
        cp_async16(kd, kbase + o);
        cp_async16(vd, vbase + o);
      } else {                                          // past the range: finite values, their probabilities are zero
        *reinterpret_cast<uint4*>(kd) = make_uint4(0u, 0u, 0u, 0u);
        *reinterpret_cast<uint4*>(vd) = make_uint4(0u, 0u, 0u, 0u);
      }
    }
    asm volatile("cp.async.commit_group;" ::: "memory");
  };
  stage(0, 0);
  // the tile's queries (already rotated by the append pass), rounded to bf16; rows past the chunk are zero
  for (int idx = threadIdx.x; idx < 64 * 32; idx += 128) {
    const int t = idx >> 5, c = idx & 31;
    uint32_t v = 0u;
    if (t < nq) {
      const float2 q = load2(qkv + (long long)(r0 + t0 + t) * 3 * d_model + h * 64 + 2 * c);
      v = pack_bf16(q.x, q.y);
    }
    reinterpret_cast<uint32_t*>(Qs + t * PM_LD)[c] = v;
  }
  __syncthreads();
  uint32_t qa[4][4];
#pragma unroll
  for (int kk = 0; kk < 4; ++kk) ldsm_x4(qa[kk], Qs + (16 * w + (lane & 15)) * PM_LD + kk * 16 + (lane >> 4) * 8);
  const int gq = lane >> 2, qq = lane & 3;
  const int pq0 = off + t0 + 16 * w + gq, pq1 = pq0 + 8;     // positions of the two query rows this thread holds
  const int warp_last = off + t0 + 16 * w + 15;              // blocks past it are skipped by this warp
  float o[8][4];
#pragma unroll
  for (int nt = 0; nt < 8; ++nt) o[nt][0] = o[nt][1] = o[nt][2] = o[nt][3] = 0.0f;
  float m_run[2] = {-INFINITY, -INFINITY}, l_run[2] = {0.0f, 0.0f};

  for (int kb = 0; kb < n_kb; ++kb) {
    const int buf = kb & 1;
    if (kb + 1 < n_kb) {
      stage(kb + 1, buf ^ 1);
      asm volatile("cp.async.wait_group 1;" ::: "memory");
    } else {
      asm volatile("cp.async.wait_group 0;" ::: "memory");
    }
    __syncthreads();
    if (kb * 64 <= warp_last) {
      const bf16* Ks = Kb[buf];
      const bf16* Vs = Vb[buf];
      float s[8][4];
#pragma unroll
      for (int nt = 0; nt < 8; ++nt) s[nt][0] = s[nt][1] = s[nt][2] = s[nt][3] = 0.0f;
#pragma unroll
      for (int kk = 0; kk < 4; ++kk) {
#pragma unroll
        for (int np = 0; np < 4; ++np) {
          uint32_t b[4];
          const int key = np * 16 + ((lane >> 4) << 3) + (lane & 7);
          ldsm_x4(b, Ks + key * PM_LD + kk * 16 + ((lane >> 3) & 1) * 8);
          mma_bf16_16816(s[2 * np], qa[kk], b[0], b[1]);
          mma_bf16_16816(s[2 * np + 1], qa[kk], b[2], b[3]);
        }
      }
      float mx[2] = {-INFINITY, -INFINITY};
#pragma unroll
      for (int nt = 0; nt < 8; ++nt) {
#pragma unroll
        for (int e = 0; e < 4; ++e) {
          const int pj = kb * 64 + nt * 8 + 2 * qq + (e & 1);
          const int pq = (e < 2) ? pq0 : pq1;
          const float v = pj <= pq ? s[nt][e] * 0.125f : -INFINITY;    // causal (transformer.py:22-29, no window)
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
          ldsm_x4_t(b, Vs + key * PM_LD + np * 16 + (lane >> 4) * 8);
          mma_bf16_16816(o[2 * np], pa[kk], b[0], b[1]);
          mma_bf16_16816(o[2 * np + 1], pa[kk], b[2], b[3]);
        }
      }
    }
    __syncthreads();   // the other buffer is overwritten by the next iteration's stage
  }
#pragma unroll
  for (int r = 0; r < 2; ++r) {
    l_run[r] += __shfl_xor_sync(0xffffffffu, l_run[r], 1);
    l_run[r] += __shfl_xor_sync(0xffffffffu, l_run[r], 2);
  }
  const int ta = t0 + 16 * w + gq, tb = ta + 8;
  if (ta < T) {
    const float inv = 1.0f / l_run[0];
    bf16* dst = out + (long long)(r0 + ta) * d_model + h * 64 + 2 * qq;
#pragma unroll
    for (int nt = 0; nt < 8; ++nt) *reinterpret_cast<uint32_t*>(dst + nt * 8) = pack_bf16(o[nt][0] * inv, o[nt][1] * inv);
  }
  if (tb < T) {
    const float inv = 1.0f / l_run[1];
    bf16* dst = out + (long long)(r0 + tb) * d_model + h * 64 + 2 * qq;
#pragma unroll
    for (int nt = 0; nt < 8; ++nt) *reinterpret_cast<uint32_t*>(dst + nt * 8) = pack_bf16(o[nt][2] * inv, o[nt][3] * inv);
  }
}

}  // namespace ptts
