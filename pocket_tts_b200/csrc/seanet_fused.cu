// Fused last SEANet decoder stage (bf16 mode): ConvTranspose1d(128 -> 64, K = 8, S = 4) -> ELU -> ResBlock
// [Conv1d(64 -> 32, K3) -> ELU -> Conv1d(32 -> 64, K1) + skip] -> ELU -> Conv1d(64 -> 1, K3) as ONE persistent tcgen05
// kernel (modules/seanet.py:20-41,141-172, modules/conv.py:93-115,151-163).  The unfused form moved the
// [streams][1920 * frames][64] activation tensors through L2 / HBM five times per pass (~570 MB at 64 streams x 4
// frames, 341 us); here a pass reads the stage input once (31 MB) and writes PCM (2 MB).
//
// One CTA per SM walks tiles of 128 TMEM lanes = input columns tau0-1 .. tau0+126 of one stream (lane 0 is a halo lane
// that only feeds its neighbour; 127 new columns per tile).  Per tile, with r = 4 output phases rho per input column:
//   GEMM-A  D_A[tau][rho*64+c]  = sum_k W_tr[rho*64+c][k] * (x[tau-1] | x[tau])[k]      M128 N256 K256, operands by TMA
//   epi-A   e = bf16(ELU(D_A + b_tr)) written by the owning thread into shared memory as the K-major SWIZZLE_128B
//           operand of the next GEMM: k-block P_rho[tau] = e(tau, rho); phases 2, 3 are written a second time one row
//           down (P2', P3': row tau+1 holds e(tau, 2..3)), which is the two-column causal history of phases 0, 1
//   GEMM-B  D_B[tau][rho*32+h]  = W3 . (P_{rho-2}, P_{rho-1}, P_rho)    (indices < 0 -> the shifted copies)   N32 K192
//   epi-B   hid = bf16(ELU(D_B + b3)) -> shared-memory operand H
//   GEMM-C  D_A[tau][rho*64+c] += W1 . hid(tau, rho)     accumulates the 1x1 conv ONTO the fp32 skip path in TMEM
//   epi-C   out = D_A + b_tr + b1;  o = bf16(ELU(out));  z_k = w_last[k] . o  (three 64-wide dots per output column)
//   pcm[4 tau + rho] = b_last + z_0(t-2) + z_1(t-1) + z_2(t)     neighbours through shared memory
// Streaming state (conv3 history = last two e columns, last conv history = last two o columns) is read from / written
// to the same history rows of b_res / b_last that hist_gather / hist_scatter move, so the engine's state handling and
// the unfused path (fp32 mode, other shapes) are unchanged.  Every output depends on its inputs through a fixed
// arithmetic order that does not depend on the tiling, so passes of 1, 2 or 4 frames stay bit-identical.
#include <cuda.h>

#include "gemm.cuh"
#include "kernels.cuh"
#include "tc_ptx.cuh"

namespace ptts {

namespace {
constexpr int FS_THREADS = 320;                       // warp 0 TMA producer, warp 1 MMA issuer, warps 2..9 epilogue
constexpr int FS_NEW = 127;                           // new input columns per tile
constexpr uint32_t FS_A_BYTES = 128 * 128;            // A k-block: 128 rows x 64 bf16
constexpr uint32_t FS_W_BYTES = 256 * 128;            // W_tr k-block: 256 rows x 64 bf16
constexpr uint32_t FS_STAGE = FS_A_BYTES + FS_W_BYTES;
constexpr uint32_t FS_OFF_E = 2 * FS_STAGE;           // 4 operand slots of 16 KB
constexpr uint32_t FS_OFF_H = FS_OFF_E + 4 * 16384;   // 2 hidden operand buffers of 16 KB (64 of 128 bytes per row used)
constexpr uint32_t FS_OFF_W3 = FS_OFF_H + 2 * 16384;  // 3 k-blocks of [32 rows][128 B]
constexpr uint32_t FS_OFF_W1 = FS_OFF_W3 + 3 * 4096;  // [64 rows][128 B]
constexpr uint32_t FS_OFF_Z = FS_OFF_W1 + 8192;       // float Z[128][4][3][2]
constexpr uint32_t FS_SMEM = FS_OFF_Z + 128 * 24 * 4 + 1024;
constexpr uint32_t TM_A = 0, TM_B = 256;              // TMEM column offsets
}  // namespace

struct FusedStage2Params {
  int n, T, tiles_per_stream, n_tiles;
  // per-channel vectors by value: kernel parameters live in the constant bank, so FADD / FFMA take them as operands
  // (no load instruction per element in the epilogues)
  float b_tr[256];     // transposed-conv bias per (phase, channel)
  float b_out[256];    // b_tr + ResBlock 1x1 bias
  float b_c3[32];
  float w_last[192];   // [tap][channel]
  float b_last;
  unsigned long long* trace;   // developer: globaltimer stamps of CTA 0, 12 per tile (nullptr: off)
  bf16* res_buf;    // [n][2 + 4T][64]: rows 0..1 conv3 history in, rows 4T..4T+1 history out
  bf16* last_buf;   // [n][2 + 4T][64]: same for the last conv
  float* pcm;       // [n][4T]
};

__device__ __forceinline__ uint32_t sw128_off(int row, int chunk) { return (uint32_t)(row * 128 + ((chunk ^ (row & 7)) << 4)); }
__device__ __forceinline__ uint32_t pack2(float a, float b) {
  const __nv_bfloat162 v = __floats2bfloat162_rn(a, b);
  return *reinterpret_cast<const uint32_t*>(&v);
}
__device__ __forceinline__ float bf16_round(float x) { return __bfloat162float(__float2bfloat16_rn(x)); }
#define FS_TRACE(slot) do { if (P.trace && blockIdx.x == 0 && lane == 0 && (warp == 1 || warp == 2) && it < 16) P.trace[it * 12 + (slot)] = tl_now(); } while (0)
__device__ __forceinline__ void fence_async_smem() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }

__device__ __forceinline__ uint4 pack8(const float* v) {
  return make_uint4(pack2(v[0], v[1]), pack2(v[2], v[3]), pack2(v[4], v[5]), pack2(v[6], v[7]));
}

// Epilogue warps (8): TMEM lane quarter q = warp & 3, channel half H (compile-time so that the per-channel constants
// are immediate constant-bank operands).
template <int H>
__device__ __forceinline__ void fused_epilogue(const FusedStage2Params& P, uint8_t* sm, uint32_t tmem_base, int warp, int lane,
                                               uint64_t* da_full, uint64_t* tmem_free, uint64_t* e_ready, uint64_t* b_full,
                                               uint64_t* h_ready, uint64_t* c_full) {
  const int q = warp & 3;
  const int r = q * 32 + lane;                            // TMEM lane = tile row
  const uint32_t t_lane = tmem_base + ((uint32_t)(q * 32) << 16);
  float* Z = reinterpret_cast<float*>(sm + FS_OFF_Z);
  const int T = P.T;
  uint32_t it = 0;
  for (int tile = blockIdx.x; tile < P.n_tiles; tile += gridDim.x, ++it) {
    const uint32_t par = it & 1u;
    const int i = tile / P.tiles_per_stream, k = tile - i * P.tiles_per_stream;
    const int tau = k * FS_NEW - 1 + r;
    const bool first = k == 0, last_col = tau == T - 1;
    bf16* res_i = P.res_buf + (size_t)i * (2 + 4 * T) * 64;
    bf16* last_i = P.last_buf + (size_t)i * (2 + 4 * T) * 64;
    uint4 keep[2][4];                                     // packed e of phases 2, 3 for the shifted copies
    // ---- epi-A pass 1: P_0..P_3
    mbar_wait(da_full, par);
    FS_TRACE(2);
    tc_fence_after();
#pragma unroll
    for (int rho = 0; rho < 4; ++rho) {
      float v[32];
      tmem_ld<32>(t_lane + TM_A + rho * 64 + H * 32, v);
#pragma unroll
      for (int j = 0; j < 32; ++j) v[j] = elu_call(v[j] + P.b_tr[rho * 64 + H * 32 + j]);
      uint8_t* slot = sm + FS_OFF_E + rho * 16384;
#pragma unroll
      for (int c = 0; c < 4; ++c) {
        const uint4 pk = pack8(v + 8 * c);
        *reinterpret_cast<uint4*>(slot + sw128_off(r, H * 4 + c)) = pk;
        if (rho >= 2) keep[rho - 2][c] = pk;
      }
      if (rho >= 2 && last_col) {                         // conv3 history out: e at output times 4T-2, 4T-1
        uint4* dst = reinterpret_cast<uint4*>(res_i + (size_t)(4 * T + rho - 2) * 64 + H * 32);
#pragma unroll
        for (int c = 0; c < 4; ++c) dst[c] = keep[rho - 2][c];
      }
    }
    fence_async_smem();
    tc_fence_before();
    __syncwarp();
    if (lane == 0) mbar_arrive_a(smem_u32(&e_ready[0]));
    FS_TRACE(3);
    // ---- epi-A pass 2: once GEMM-B of phases 2, 3 has consumed P_2, P_3, their slots take the copies shifted down
    // one row (row tau+1 <- e(tau, 2..3)); the stream's first column takes the carried history instead
    mbar_wait(&b_full[3], par);
    FS_TRACE(4);
    if (first && r == 0) {
#pragma unroll
      for (int p = 0; p < 2; ++p) {
        const uint4* src = reinterpret_cast<const uint4*>(res_i + (size_t)p * 64 + H * 32);
#pragma unroll
        for (int c = 0; c < 4; ++c) keep[p][c] = src[c];
      }
    }
    if (r < 127) {
#pragma unroll
      for (int p = 0; p < 2; ++p) {
        uint8_t* slot = sm + FS_OFF_E + (2 + p) * 16384;
#pragma unroll
        for (int c = 0; c < 4; ++c) *reinterpret_cast<uint4*>(slot + sw128_off(r + 1, H * 4 + c)) = keep[p][c];
      }
    }
    fence_async_smem();
    __syncwarp();
    if (lane == 0) mbar_arrive_a(smem_u32(&e_ready[1]));
    FS_TRACE(5);
    // ---- epi-B: hid = ELU(D_B + b3) -> H[rho & 1]; order 2, 3, 0, 1 (H[rho & 1] is free once GEMM-C(rho ^ 2) is done)
#pragma unroll
    for (int o4 = 0; o4 < 4; ++o4) {
      const int rho = (o4 + 2) & 3;
      mbar_wait(&b_full[rho], par);
      tc_fence_after();
      float v[16];
      tmem_ld<16>(t_lane + TM_B + rho * 32 + H * 16, v);
      if (o4 >= 2) mbar_wait(&c_full[rho + 2], par);
#pragma unroll
      for (int j = 0; j < 16; ++j) v[j] = elu_call(v[j] + P.b_c3[H * 16 + j]);
      uint8_t* hb = sm + FS_OFF_H + (rho & 1) * 16384;
#pragma unroll
      for (int c = 0; c < 2; ++c) *reinterpret_cast<uint4*>(hb + sw128_off(r, H * 2 + c)) = pack8(v + 8 * c);
      fence_async_smem();
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive_a(smem_u32(&h_ready[rho]));
    }
    FS_TRACE(6);
    // ---- epi-C: out = D_A + b_tr + b1, o = bf16(ELU(out)), last-conv partial dots
#pragma unroll
    for (int o4 = 0; o4 < 4; ++o4) {
      const int rho = (o4 + 2) & 3;
      mbar_wait(&c_full[rho], par);
      tc_fence_after();
      float v[32];
      tmem_ld<32>(t_lane + TM_A + rho * 64 + H * 32, v);
#pragma unroll
      for (int j = 0; j < 32; ++j) v[j] = bf16_round(elu_call(v[j] + P.b_out[rho * 64 + H * 32 + j]));
      if (rho >= 2) {
        if (first && r == 0) {                            // last-conv history in: o at output times -2, -1
          const bf16* src = last_i + (size_t)(rho - 2) * 64 + H * 32;
#pragma unroll
          for (int j = 0; j < 32; ++j) v[j] = __bfloat162float(src[j]);
        }
        if (last_col) {                                   // history out: o at output times 4T-2, 4T-1
          uint4* dst = reinterpret_cast<uint4*>(last_i + (size_t)(4 * T + rho - 2) * 64 + H * 32);
#pragma unroll
          for (int c = 0; c < 4; ++c) dst[c] = pack8(v + 8 * c);
        }
      }
      float z0 = 0.f, z1 = 0.f, z2 = 0.f;
#pragma unroll
      for (int j = 0; j < 32; ++j) {
        z0 = fmaf(P.w_last[H * 32 + j], v[j], z0);
        z1 = fmaf(P.w_last[64 + H * 32 + j], v[j], z1);
        z2 = fmaf(P.w_last[128 + H * 32 + j], v[j], z2);
      }
      float* zr = Z + ((r * 4 + rho) * 3) * 2 + H;
      zr[0] = z0; zr[2] = z1; zr[4] = z2;
    }
    tc_fence_before();
    __syncwarp();
    if (lane == 0) mbar_arrive_a(smem_u32(tmem_free));    // TMEM drained: the next tile's GEMM-A may start
    FS_TRACE(7);
    asm volatile("bar.sync 1, 256;" ::: "memory");        // Z complete
    if (r >= 1 && tau >= 0 && tau < T) {
      auto zz = [&](int rr, int rho, int kk) { const float* p = Z + ((rr * 4 + rho) * 3 + kk) * 2; return p[0] + p[1]; };
      float2 out;
      if (H == 0) {
        out.x = P.b_last + zz(r - 1, 2, 0) + zz(r - 1, 3, 1) + zz(r, 0, 2);
        out.y = P.b_last + zz(r - 1, 3, 0) + zz(r, 0, 1) + zz(r, 1, 2);
      } else {
        out.x = P.b_last + zz(r, 0, 0) + zz(r, 1, 1) + zz(r, 2, 2);
        out.y = P.b_last + zz(r, 1, 0) + zz(r, 2, 1) + zz(r, 3, 2);
      }
      *reinterpret_cast<float2*>(P.pcm + (size_t)i * 4 * T + 4 * (size_t)tau + 2 * H) = out;
    }
    asm volatile("bar.sync 2, 256;" ::: "memory");        // Z may be overwritten by the next tile
    FS_TRACE(8);
  }
}

__global__ void __launch_bounds__(FS_THREADS, 1)
seanet_stage2_fused_kernel(const __grid_constant__ CUtensorMap tm_x, const __grid_constant__ CUtensorMap tm_wtr,
                           const __grid_constant__ CUtensorMap tm_w3, const __grid_constant__ CUtensorMap tm_w1,
                           const __grid_constant__ FusedStage2Params P) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* sm = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  __shared__ __align__(8) uint64_t full_bar[2], empty_bar[2];
  __shared__ __align__(8) uint64_t w_bar, da_full, tmem_free, e_ready[2], b_full[4], h_ready[4], c_full[4];
  __shared__ uint32_t tmem_base_smem;

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  pdl_launch_dependents();
  if (warp == 0 && lane == 0) {
    asm volatile("prefetch.tensormap [%0];" ::"l"(&tm_x) : "memory");
    asm volatile("prefetch.tensormap [%0];" ::"l"(&tm_wtr) : "memory");
    for (int s = 0; s < 2; ++s) { mbar_init(&full_bar[s], 1); mbar_init(&empty_bar[s], 1); mbar_init(&e_ready[s], 8); }
    mbar_init(&w_bar, 1); mbar_init(&da_full, 1); mbar_init(&tmem_free, 8);
    for (int r = 0; r < 4; ++r) { mbar_init(&b_full[r], 1); mbar_init(&h_ready[r], 8); mbar_init(&c_full[r], 1); }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    // resident ResBlock weights (depend on nothing): W3 as three k-blocks of 32 rows, W1 as one block of 64 rows
    mbar_expect_tx(&w_bar, 3 * 4096 + 8192);
    for (int j = 0; j < 3; ++j) tma_load_2d(sm + FS_OFF_W3 + j * 4096, &tm_w3, &w_bar, j * 64, 0);
    tma_load_2d(sm + FS_OFF_W1, &tm_w1, &w_bar, 0, 0);
  }
  if (warp == 1) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&tmem_base_smem)), "n"(512) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  pdl_wait();
  const uint32_t tmem_base = tmem_base_smem;
  const uint32_t sm_a = smem_u32(sm);
  const uint32_t full_a = smem_u32(&full_bar[0]), empty_a = smem_u32(&empty_bar[0]);

  if (warp == 0) {
    // ===== TMA producer: 4 k-blocks of (A 128 x 64, W_tr 256 x 64) per tile through a 2-stage ring =====
    uint32_t s = 0, ph = 0;
    for (int tile = blockIdx.x; tile < P.n_tiles; tile += gridDim.x) {
      const int i = tile / P.tiles_per_stream, k = tile - i * P.tiles_per_stream;
      const int row0 = i * (1 + P.T) + k * FS_NEW - 1;     // A row tau covers x[tau-1] | x[tau]; may be -1 (zero fill)
      for (int kb = 0; kb < 4; ++kb) {
        mbar_wait_a(empty_a + 8u * s, ph ^ 1u);
        if (lane == 0) {
          const uint32_t st = sm_a + s * FS_STAGE, fb = full_a + 8u * s;
          mbar_expect_tx_a(fb, FS_STAGE);
          tma_load_2d_a(st, &tm_x, fb, kb * 64, row0);
          tma_load_2d_a(st + FS_A_BYTES, &tm_wtr, fb, kb * 64, 0);
        }
        __syncwarp();
        if (++s == 2) { s = 0; ph ^= 1u; }
      }
    }
  } else if (warp == 1) {
    // ===== MMA issuer =====
    const uint32_t idesc_a = (1u << 4) | (1u << 7) | (1u << 10) | ((256u >> 3) << 17) | ((128u >> 4) << 24);
    const uint32_t idesc_b = (1u << 4) | (1u << 7) | (1u << 10) | ((32u >> 3) << 17) | ((128u >> 4) << 24);
    const uint32_t idesc_c = (1u << 4) | (1u << 7) | (1u << 10) | ((64u >> 3) << 17) | ((128u >> 4) << 24);
    const uint64_t d_ring = make_sw128_desc(sm_a), d_e = make_sw128_desc(sm_a + FS_OFF_E), d_h = make_sw128_desc(sm_a + FS_OFF_H);
    const uint64_t d_w3 = make_sw128_desc(sm_a + FS_OFF_W3), d_w1 = make_sw128_desc(sm_a + FS_OFF_W1);
    uint32_t s = 0, ph = 0, it = 0;
    mbar_wait(&w_bar, 0);
    for (int tile = blockIdx.x; tile < P.n_tiles; tile += gridDim.x, ++it) {
      const uint32_t par = it & 1u;
      mbar_wait(&tmem_free, par ^ 1u);            // epilogue of the previous tile has drained TMEM (first tile: passes)
      FS_TRACE(0);
      tc_fence_after();
      for (int kb = 0; kb < 4; ++kb) {            // GEMM-A
        mbar_wait_a(full_a + 8u * s, ph);
        tc_fence_after();
        if (lane == 0) {
          const uint64_t soff = (uint64_t)((s * FS_STAGE) >> 4);
          const uint64_t ad = d_ring + soff, bd = d_ring + soff + (uint64_t)(FS_A_BYTES >> 4);
#pragma unroll
          for (int k = 0; k < 4; ++k) tc_mma_bf16(tmem_base + TM_A, ad + (uint64_t)(k * 2), bd + (uint64_t)(k * 2), idesc_a, (kb | k) != 0);
          tc_commit_a(empty_a + 8u * s);
          if (kb == 3) tc_commit(&da_full);
        }
        __syncwarp();
        if (++s == 2) { s = 0; ph ^= 1u; }
      }
      auto gemm_b = [&](int rho) {                // window (P_{rho-2}, P_{rho-1}, P_rho) = slots (rho-2+j) & 3
        if (lane == 0) {
#pragma unroll
          for (int j = 0; j < 3; ++j) {
            const uint64_t ad = d_e + (uint64_t)((((rho + 2 + j) & 3) * 16384) >> 4), bd = d_w3 + (uint64_t)((j * 4096) >> 4);
#pragma unroll
            for (int k = 0; k < 4; ++k)
              tc_mma_bf16(tmem_base + TM_B + rho * 32, ad + (uint64_t)(k * 2), bd + (uint64_t)(k * 2), idesc_b, (j | k) != 0);
          }
          tc_commit(&b_full[rho]);
        }
        __syncwarp();
      };
      auto gemm_c = [&](int rho) {                // D_A[:, rho*64 ..] += hid(rho) W1^T, K = 32
        mbar_wait(&h_ready[rho], par);
        tc_fence_after();
        if (lane == 0) {
          const uint64_t ad = d_h + (uint64_t)(((rho & 1) * 16384) >> 4);
#pragma unroll
          for (int k = 0; k < 2; ++k)
            tc_mma_bf16(tmem_base + TM_A + rho * 64, ad + (uint64_t)(k * 2), d_w1 + (uint64_t)(k * 2), idesc_c, 1u);
          tc_commit(&c_full[rho]);
        }
        __syncwarp();
      };
      FS_TRACE(1);
      mbar_wait(&e_ready[0], par);
      tc_fence_after();
      gemm_b(2); gemm_b(3);
      FS_TRACE(9);
      mbar_wait(&e_ready[1], par);
      tc_fence_after();
      gemm_b(0); gemm_b(1);
      gemm_c(2); gemm_c(3); gemm_c(0); gemm_c(1);
      FS_TRACE(10);
    }
  } else {
    // ===== epilogue warps 2..9: TMEM lane quarter q, channel half h =====
    if (((warp - 2) >> 2) == 0) fused_epilogue<0>(P, sm, tmem_base, warp, lane, &da_full, &tmem_free, e_ready, b_full, h_ready, c_full);
    else fused_epilogue<1>(P, sm, tmem_base, warp, lane, &da_full, &tmem_free, e_ready, b_full, h_ready, c_full);
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "n"(512) : "memory");
  }
}

bool seanet_stage2_fused_supported(int c_in, int c_out, int ratio, int hid, int hid_padded, int res_kernel, int last_kernel) {
  return c_in == 128 && c_out == 64 && ratio == 4 && hid == 32 && hid_padded == 64 && res_kernel == 3 && last_kernel == 3;
}

// x_buf [n][1 + T][128] bf16 (row 0 = carried input column), w_tr [256][256], w3 [64 (32 used)][192], w1 [64][64 (32 used)]
cudaError_t launch_seanet_stage2_fused(const bf16* x_buf, const bf16* w_tr, const bf16* w3, const bf16* w1, const FusedStage2Consts& cst,
                                       bf16* res_buf, bf16* last_buf, float* pcm, int n, int T, int grid_cap, bool pdl, cudaStream_t st,
                                       unsigned long long* trace) {
  if (n <= 0 || T <= 0) return cudaSuccess;
  CUtensorMap tx, twtr, tw3, tw1;
  if (!tc_make_map(&tx, x_buf, (long long)n * (1 + T) - 1, 256, 128, 128)) return cudaErrorUnknown;
  if (!tc_make_map(&twtr, w_tr, 256, 256, 256, 256)) return cudaErrorUnknown;
  if (!tc_make_map(&tw3, w3, 64, 192, 192, 32)) return cudaErrorUnknown;
  if (!tc_make_map(&tw1, w1, 64, 64, 64, 64)) return cudaErrorUnknown;
  FusedStage2Params P{};
  P.n = n; P.T = T;
  P.tiles_per_stream = (T + FS_NEW - 1) / FS_NEW;
  P.n_tiles = n * P.tiles_per_stream;
  for (int j = 0; j < 256; ++j) { P.b_tr[j] = cst.b_tr[j]; P.b_out[j] = cst.b_tr[j] + cst.b_c1[j & 63]; }
  for (int j = 0; j < 32; ++j) P.b_c3[j] = cst.b_c3[j];
  for (int j = 0; j < 192; ++j) P.w_last[j] = cst.w_last[j];
  P.b_last = cst.b_last;
  P.trace = trace;
  P.res_buf = res_buf; P.last_buf = last_buf; P.pcm = pcm;
  static DeviceOnce attr_done;
  if (attr_done.first()) {
    cudaError_t e = cudaFuncSetAttribute(seanet_stage2_fused_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)FS_SMEM);
    if (e != cudaSuccess) return e;
  }
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = dim3(P.n_tiles < grid_cap ? P.n_tiles : grid_cap, 1, 1);
  cfg.blockDim = dim3(FS_THREADS, 1, 1);
  cfg.dynamicSmemBytes = FS_SMEM;
  cfg.stream = st;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = pdl ? 1 : 0;
  return cudaLaunchKernelEx(&cfg, seanet_stage2_fused_kernel, tx, twtr, tw3, tw1, P);
}


// =====================================================================================================================
// Fused SEANet ResBlock (stages 0 and 1 of the decoder, bf16 mode):  out = ELU( y + b1 + W1 . ELU( b3 + W3 . (e[t-2] | e[t-1] | e[t]) ) )
// with e = ELU(y) (seanet.py:20-41).  The unfused form ran conv3 and conv1 as two GEMM launches with the hidden tensor
// (padded to 64 channels) written to and read back from L2 / HBM in between.  Here one persistent CTA per SM walks
// tiles of 128 time steps: GEMM-B (K = 3C through a TMA ring, A rows overlap in the tensor map) -> D_B in TMEM ->
// epilogue writes bf16(ELU(D_B + b3)) into shared memory in the K-major SWIZZLE_128B form -> GEMM-C with the resident
// W1 -> D_C -> epilogue adds bias + skip (raw y from global) and stores ELU(.) as the next transposed conv's input.
// C = 128 (stage 1): TMEM 2 x (64 + 128) columns, two tiles in flight (GEMM-B of tile t+1 runs beside the epilogues of
// tile t); C = 256 (stage 0): 128 + 256 columns, one tile in flight.
// =====================================================================================================================
struct ResBlockParams {
  int n, T, tiles_per_stream, n_tiles;
  float b3[128];       // conv3 bias (hidden width <= 128)
  float b1[256];       // conv1 bias
  const bf16* raw;     // [n][T][C]   y (skip)
  bf16* out;           // [n][out_hist + T][C]: ELU(out) behind the next conv's history rows
  int out_hist;
};

template <int C, int NBUF>
__global__ void __launch_bounds__(FS_THREADS, 1)
seanet_resblock_fused_kernel(const __grid_constant__ CUtensorMap tm_a, const __grid_constant__ CUtensorMap tm_w3,
                             const __grid_constant__ CUtensorMap tm_w1, const __grid_constant__ ResBlockParams P) {
  constexpr int HP = C / 2;                       // hidden channels (64 / 128)
  constexpr int NKB = 3 * C / 64;                 // k-blocks of GEMM-B
  constexpr int HKB = HP / 64;                    // k-blocks of GEMM-C
  constexpr uint32_t A_BYTES = 128 * 128, W3_BYTES = HP * 128, STAGE = A_BYTES + W3_BYTES;
  constexpr int STAGES = C == 128 ? 4 : 3;
  constexpr uint32_t OFF_H = STAGES * STAGE;                         // [NBUF][HKB][128 rows][128 B]
  constexpr uint32_t OFF_W1 = OFF_H + NBUF * HKB * 16384;            // [HKB][C rows][128 B]
  constexpr uint32_t TM_DB = 0, TM_DC = HP, TM_STRIDE = HP + C;      // per buffer: D_B | D_C
  constexpr uint32_t TM_COLS = 512;
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* sm = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  __shared__ __align__(8) uint64_t full_bar[4], empty_bar[4];
  __shared__ __align__(8) uint64_t w_bar, db_full[2], h_ready[2], dc_full[2], tmem_free[2];
  __shared__ uint32_t tmem_base_smem;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  pdl_launch_dependents();
  if (warp == 0 && lane == 0) {
    asm volatile("prefetch.tensormap [%0];" ::"l"(&tm_a) : "memory");
    asm volatile("prefetch.tensormap [%0];" ::"l"(&tm_w3) : "memory");
    for (int s = 0; s < STAGES; ++s) { mbar_init(&full_bar[s], 1); mbar_init(&empty_bar[s], 1); }
    mbar_init(&w_bar, 1);
    for (int b = 0; b < 2; ++b) { mbar_init(&db_full[b], 1); mbar_init(&h_ready[b], 8); mbar_init(&dc_full[b], 1); mbar_init(&tmem_free[b], 8); }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    mbar_expect_tx(&w_bar, HKB * C * 128);
    for (int j = 0; j < HKB; ++j) tma_load_2d(sm + OFF_W1 + j * C * 128, &tm_w1, &w_bar, j * 64, 0);
  }
  if (warp == 1) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&tmem_base_smem)), "n"(TM_COLS) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  pdl_wait();
  const uint32_t tmem_base = tmem_base_smem;
  const uint32_t sm_a = smem_u32(sm);
  const uint32_t full_a = smem_u32(&full_bar[0]), empty_a = smem_u32(&empty_bar[0]);

  if (warp == 0) {
    // ===== TMA producer: NKB k-blocks of (A 128 x 64, W3 HP x 64) per tile =====
    uint32_t s = 0, ph = 0;
    for (int tile = blockIdx.x; tile < P.n_tiles; tile += gridDim.x) {
      const int i = tile / P.tiles_per_stream, k = tile - i * P.tiles_per_stream;
      const int row0 = i * (2 + P.T) + k * 128;        // A row t = buffer rows t .. t+2 = e[t-2 .. t]
      for (int kb = 0; kb < NKB; ++kb) {
        mbar_wait_a(empty_a + 8u * s, ph ^ 1u);
        if (lane == 0) {
          const uint32_t st = sm_a + s * STAGE, fb = full_a + 8u * s;
          mbar_expect_tx_a(fb, STAGE);
          tma_load_2d_a(st, &tm_a, fb, kb * 64, row0);
          tma_load_2d_a(st + A_BYTES, &tm_w3, fb, kb * 64, 0);
        }
        __syncwarp();
        if (++s == STAGES) { s = 0; ph ^= 1u; }
      }
    }
  } else if (warp == 1) {
    // ===== MMA issuer: B(0), B(1), C(0), B(2), C(1), ... with NBUF = 2; B(t), C(t) with NBUF = 1 =====
    const uint32_t idesc_b = (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(HP >> 3) << 17) | ((128u >> 4) << 24);
    const uint32_t idesc_c = (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(C >> 3) << 17) | ((128u >> 4) << 24);
    const uint64_t d_ring = make_sw128_desc(sm_a), d_h = make_sw128_desc(sm_a + OFF_H), d_w1 = make_sw128_desc(sm_a + OFF_W1);
    uint32_t s = 0, ph = 0;
    mbar_wait(&w_bar, 0);
    int n_mine = 0;
    for (int tile = blockIdx.x; tile < P.n_tiles; tile += gridDim.x) ++n_mine;
    auto gemm_b = [&](int t) {
      const uint32_t b = NBUF == 2 ? (uint32_t)(t & 1) : 0u, use = (uint32_t)(t / NBUF);
      mbar_wait(&tmem_free[b], (use & 1u) ^ 1u);       // the epilogue drained this buffer (first use: passes)
      tc_fence_after();
      for (int kb = 0; kb < NKB; ++kb) {
        mbar_wait_a(full_a + 8u * s, ph);
        tc_fence_after();
        if (lane == 0) {
          const uint64_t soff = (uint64_t)((s * STAGE) >> 4);
          const uint64_t ad = d_ring + soff, bd = d_ring + soff + (uint64_t)(A_BYTES >> 4);
#pragma unroll
          for (int k = 0; k < 4; ++k)
            tc_mma_bf16(tmem_base + b * TM_STRIDE + TM_DB, ad + (uint64_t)(k * 2), bd + (uint64_t)(k * 2), idesc_b, (kb | k) != 0);
          tc_commit_a(empty_a + 8u * s);
          if (kb == NKB - 1) tc_commit(&db_full[b]);
        }
        __syncwarp();
        if (++s == STAGES) { s = 0; ph ^= 1u; }
      }
    };
    auto gemm_c = [&](int t) {
      const uint32_t b = NBUF == 2 ? (uint32_t)(t & 1) : 0u, use = (uint32_t)(t / NBUF);
      mbar_wait(&h_ready[b], use & 1u);
      tc_fence_after();
      if (lane == 0) {
#pragma unroll
        for (int j = 0; j < HKB; ++j) {
          const uint64_t ad = d_h + (uint64_t)(((b * HKB + j) * 16384) >> 4), bd = d_w1 + (uint64_t)((j * C * 128) >> 4);
#pragma unroll
          for (int k = 0; k < 4; ++k)
            tc_mma_bf16(tmem_base + b * TM_STRIDE + TM_DC, ad + (uint64_t)(k * 2), bd + (uint64_t)(k * 2), idesc_c, (j | k) != 0);
        }
        tc_commit(&dc_full[b]);
      }
      __syncwarp();
    };
    if (NBUF == 2) {
      for (int t = 0; t < n_mine; ++t) {
        gemm_b(t);
        if (t > 0) gemm_c(t - 1);
      }
      if (n_mine > 0) gemm_c(n_mine - 1);
    } else {
      for (int t = 0; t < n_mine; ++t) { gemm_b(t); gemm_c(t); }
    }
  } else {
    // ===== epilogue warps 2..9: TMEM lane quarter q, column half hh =====
    const int q = warp & 3, hh = (warp - 2) >> 2;
    const int r = q * 32 + lane;
    const uint32_t t_lane = tmem_base + ((uint32_t)(q * 32) << 16);
    int t = 0;
    for (int tile = blockIdx.x; tile < P.n_tiles; tile += gridDim.x, ++t) {
      const uint32_t b = NBUF == 2 ? (uint32_t)(t & 1) : 0u, par = (uint32_t)(t / NBUF) & 1u;
      const int i = tile / P.tiles_per_stream, k = tile - i * P.tiles_per_stream;
      const int ts = k * 128 + r;                       // time step of this row
      const bool valid = ts < P.T;
      // ---- epi-B: hidden = bf16(ELU(D_B + b3)) -> H[b]
      mbar_wait(&db_full[b], par);
      tc_fence_after();
#pragma unroll
      for (int c0 = 0; c0 < HP / 2; c0 += 32) {
        const int col = hh * (HP / 2) + c0;
        float v[32];
        tmem_ld<32>(t_lane + b * TM_STRIDE + TM_DB + col, v);
#pragma unroll
        for (int j = 0; j < 32; ++j) v[j] = elu_call(v[j] + P.b3[col + j]);
#pragma unroll
        for (int c = 0; c < 4; ++c) {
          const int chunk = (col >> 3) + c;             // 16-byte chunk of the hidden row
          *reinterpret_cast<uint4*>(sm + OFF_H + (b * HKB + (chunk >> 3)) * 16384 + sw128_off(r, chunk & 7)) = pack8(v + 8 * c);
        }
      }
      fence_async_smem();
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive_a(smem_u32(&h_ready[b]));
      // ---- epi-C: out = ELU(D_C + b1 + y)
      const bf16* raw = P.raw + ((size_t)i * P.T + (valid ? ts : 0)) * C + hh * (C / 2);
      bf16* dst = P.out + ((size_t)i * (P.out_hist + P.T) + P.out_hist + (valid ? ts : 0)) * C + hh * (C / 2);
      mbar_wait(&dc_full[b], par);
      tc_fence_after();
#pragma unroll
      for (int c0 = 0; c0 < C / 2; c0 += 32) {
        float v[32], y[32];
        tmem_ld<32>(t_lane + b * TM_STRIDE + TM_DC + hh * (C / 2) + c0, v);
        if (valid) {
          row_load<32>(raw + c0, y);
#pragma unroll
          for (int j = 0; j < 32; ++j) v[j] = elu_call(v[j] + P.b1[hh * (C / 2) + c0 + j] + y[j]);
          row_store<32>(dst + c0, v);
        }
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive_a(smem_u32(&tmem_free[b]));
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "n"(TM_COLS) : "memory");
  }
}

bool seanet_resblock_fused_supported(int C, int hid, int hid_padded, int res_kernel) {
  return (C == 128 || C == 256) && hid * 2 == C && hid_padded == hid && res_kernel == 3;
}

// e_buf [n][2 + T][C] (ELU(y) behind two history rows), raw [n][T][C], w3 [C/2][3C], w1 [C][C/2]; out [n][out_hist + T][C]
template <int C, int NBUF>
static cudaError_t launch_resblock_t(const bf16* e_buf, const bf16* raw, const bf16* w3, const bf16* w1, const float* b3, const float* b1,
                                     bf16* out, int out_hist, int n, int T, int n_sms, bool pdl, cudaStream_t st) {
  constexpr int HP = C / 2;
  CUtensorMap ta, tw3, tw1;
  if (!tc_make_map(&ta, e_buf, (long long)n * (2 + T) - 2, 3 * C, C, 128)) return cudaErrorUnknown;
  if (!tc_make_map(&tw3, w3, HP, 3 * C, 3 * C, HP)) return cudaErrorUnknown;
  if (!tc_make_map(&tw1, w1, C, HP, HP, C)) return cudaErrorUnknown;
  ResBlockParams P{};
  P.n = n; P.T = T;
  P.tiles_per_stream = (T + 127) / 128;
  P.n_tiles = n * P.tiles_per_stream;
  for (int j = 0; j < HP; ++j) P.b3[j] = b3[j];
  for (int j = 0; j < C; ++j) P.b1[j] = b1[j];
  P.raw = raw; P.out = out; P.out_hist = out_hist;
  constexpr int STAGES = C == 128 ? 4 : 3, HKB = HP / 64;
  const size_t smem = (size_t)STAGES * (128 * 128 + HP * 128) + (size_t)NBUF * HKB * 16384 + (size_t)HKB * C * 128 + 1024;
  auto kern = seanet_resblock_fused_kernel<C, NBUF>;
  static DeviceOnce attr_done;
  if (attr_done.first()) {
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
  }
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = dim3(P.n_tiles < n_sms ? P.n_tiles : n_sms, 1, 1);
  cfg.blockDim = dim3(FS_THREADS, 1, 1);
  cfg.dynamicSmemBytes = smem;
  cfg.stream = st;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = pdl ? 1 : 0;
  return cudaLaunchKernelEx(&cfg, kern, ta, tw3, tw1, P);
}

cudaError_t launch_seanet_resblock_fused(int C, const bf16* e_buf, const bf16* raw, const bf16* w3, const bf16* w1, const float* b3_host,
                                         const float* b1_host, bf16* out, int out_hist, int n, int T, int n_sms, bool pdl, cudaStream_t st) {
  if (n <= 0 || T <= 0) return cudaSuccess;
  if (C == 128) return launch_resblock_t<128, 2>(e_buf, raw, w3, w1, b3_host, b1_host, out, out_hist, n, T, n_sms, pdl, st);
  if (C == 256) return launch_resblock_t<256, 1>(e_buf, raw, w3, w1, b3_host, b1_host, out, out_hist, n, T, n_sms, pdl, st);
  return cudaErrorInvalidValue;
}

}  // namespace ptts
