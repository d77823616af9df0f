// Backward of the non-causal head-dim-64 attention on tcgen05 tensor cores (training step, BASELINE.json configs[4]).
//
// What autograd does behind F.scaled_dot_product_attention (reference src/f5_tts/model/modules.py:499-508) when
// trainer.py:384 calls backward:   P = softmax(scale Q K^T),  dV = P^T dO,  dP = dO V^T,  dS = P o (dP - D),
// D = rowsum(dO o O),  dQ = scale dS K,  dK = scale dS^T Q.  P is never stored: the forward kernel (tc_attn.cu) hands
// over the per-row log-sum-exp and both kernels here recompute P = exp2(scale_log2 S - lse2) tile by tile.
//
// Two kernels, both deterministic (every output element has ONE owner CTA, accumulation order fixed):
//   tc_attn_dq_kernel  : a CTA owns 128 query rows of one (head, batch element) and walks the key blocks;
//                        S = Q K_j^T and dP = dO V_j^T into TMEM, thread r turns its row into dS (bf16, back into
//                        TMEM), dQ += dS K_j with dS as the TMEM A operand and K_j read MN-major.
//   tc_attn_dkv_kernel : a CTA owns 128 keys and walks the query blocks; the TRANSPOSED products S^T = K Q_i^T and
//                        dP^T = V dO_i^T put the key on the TMEM lane, so P^T and dS^T are again TMEM A operands:
//                        dV += P^T dO_i, dK += dS^T Q_i (Q_i, dO_i read MN-major).  Per-query lse2 / D come from smem.
// 4 compute warps (thread = TMEM lane), one TMA warp, one MMA warp; operands are the forward's 128 x 64 bf16 tiles
// (128-byte swizzle), so every tile serves as a K-major operand of one product and an MN-major operand of another.
// Inside a CTA the MMAs of a step and its exponentials do not overlap (one S / dP buffer); instead a step is 64 columns
// wide, which fits TWO CTAs on an SM (tensor memory 192 / 256 columns, 80 KB of shared memory, <= 168 registers), and
// one CTA's MMAs run under the other's exponentials.
#include "common.cuh"
#include "sm100_ptx.cuh"
#include "tc_gemm.h"

#include <cstdlib>

namespace f5b {

using namespace ptx;

static constexpr int AB_D = 64;
static constexpr int AB_TILE = 128 * 64 * 2;  // one 128-row x 64-column bf16 tile: 16 KB
static constexpr int AB_STAGES = 3;
// KB = columns of S per step (keys in the dQ kernel, queries in the dK/dV kernel).  KB = 64: 192 (dQ) / 256 (dK/dV)
// tensor-memory columns and 80 KB of shared memory per CTA, so TWO CTAs share an SM and one's MMAs run under the other's
// exponentials (inside a CTA the two do not overlap: one S / dP buffer).  KB = 128: one CTA per SM (comparison).
template <int KB>
struct AbCfg {
  static constexpr int STREAM_TILE = KB * 64 * 2;  // one streamed tile: KB rows x 64 columns of bf16
  static constexpr int SMEM = 2 * AB_TILE + 2 * AB_STAGES * STREAM_TILE + 1024 /* alignment */ + 256 /* barriers */ + 2 * 2 * KB * 4;
  static constexpr uint32_t TMEM_COLS = KB == 64 ? 256 : 512;
};

__device__ __forceinline__ float ab_exp2(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}

// ------------------------------------------------------------------------------------------------------------------
// dQ: grid (query tiles, heads, batches).  TMEM: S [0,KB) (dS bf16 pairs over [0,KB/2)), dP [KB,2KB), dQ [2KB,2KB+64)
// ------------------------------------------------------------------------------------------------------------------
template <int KB>
__global__ void __launch_bounds__(192, KB == 64 ? 2 : 1)
tc_attn_dq_kernel(const __grid_constant__ CUtensorMap tm_qkv, const __grid_constant__ CUtensorMap tm_kv,
                  const __grid_constant__ CUtensorMap tm_do,
                  const float* __restrict__ lse2, const float* __restrict__ Dr, float* __restrict__ dqkv, int rows,
                  int heads, const int* __restrict__ kv_len, float scale, float scale_log2) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
  uint8_t* sQ = smem;
  uint8_t* sDO = sQ + AB_TILE;
  constexpr int ST = AbCfg<KB>::STREAM_TILE;
  constexpr uint32_t C_DP = KB, C_DQ = 2 * KB;  // TMEM columns: S [0, KB), dP [KB, 2 KB), dQ [2 KB, 2 KB + 64)
  uint8_t* sK = sDO + AB_TILE;       // AB_STAGES
  uint8_t* sV = sK + AB_STAGES * ST;  // AB_STAGES
  uint64_t* bars = reinterpret_cast<uint64_t*>(sV + AB_STAGES * ST);
  uint64_t* q_full = bars;                   // [1]
  uint64_t* kv_full = bars + 1;              // [AB_STAGES]
  uint64_t* kv_empty = kv_full + AB_STAGES;  // [AB_STAGES]
  uint64_t* s_full = kv_empty + AB_STAGES;   // MMA -> compute: S and dP of block j are in TMEM
  uint64_t* p_full = s_full + 1;             // compute -> MMA: dS of block j is in TMEM
  uint64_t* o_done = p_full + 1;             // MMA -> compute: dQ is complete
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(o_done + 1);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int q0 = blockIdx.x * 128, h = blockIdx.y, b = blockIdx.z;
  const int inner = heads * AB_D;
  const int L = kv_len ? kv_len[b] : rows;
  const int nkv = (L + KB - 1) / KB;

  if (threadIdx.x == 0) {
    prefetch_tmap(&tm_qkv);
    prefetch_tmap(&tm_do);
    prefetch_tmap(&tm_kv);
    mbar_init(q_full, 1);
    for (int i = 0; i < AB_STAGES; ++i) {
      mbar_init(&kv_full[i], 1);
      mbar_init(&kv_empty[i], 1);
    }
    mbar_init(s_full, 1);
    mbar_init(p_full, 4);
    mbar_init(o_done, 1);
    fence_barrier_init();
  }
  if (warp == 5) tmem_alloc(tmem_slot, AbCfg<KB>::TMEM_COLS);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 4) {
    if (elect_one()) {
      mbar_expect_tx(q_full, 2 * AB_TILE);
      tma_load_3d(sQ, &tm_qkv, q_full, h * AB_D, q0, b);
      tma_load_3d(sDO, &tm_do, q_full, h * AB_D, q0, b);
      for (int j = 0; j < nkv; ++j) {
        const int st = j % AB_STAGES;
        mbar_wait(&kv_empty[st], ((j / AB_STAGES) & 1) ^ 1);
        mbar_expect_tx(&kv_full[st], 2 * ST);
        tma_load_3d(sK + st * ST, &tm_kv, &kv_full[st], inner + h * AB_D, j * KB, b);
        tma_load_3d(sV + st * ST, &tm_kv, &kv_full[st], 2 * inner + h * AB_D, j * KB, b);
      }
    }
  } else if (warp == 5) {
    if (elect_one()) {
      constexpr uint32_t IDESC_S = umma_idesc_bf16(128, KB, 0, 0);
      constexpr uint32_t IDESC_O = umma_idesc_bf16(128, 64, 0, 1);  // B read MN-major
      const uint64_t dKm = umma_desc_sw128(0, 16, 1024);    // K-major view of a tile
      const uint64_t dMN = umma_desc_sw128(0, 1024, 1024);  // MN-major view of the same tile
      const uint32_t q_lo = smem_u32(sQ) >> 4, do_lo = smem_u32(sDO) >> 4, k_lo = smem_u32(sK) >> 4,
                     v_lo = smem_u32(sV) >> 4;
      mbar_wait(q_full, 0);
      for (int j = 0; j < nkv; ++j) {
        const int st = j % AB_STAGES;
        mbar_wait(&kv_full[st], (j / AB_STAGES) & 1);
        tc_fence_after();
        const uint64_t qd = dKm | (uint64_t)q_lo, od = dKm | (uint64_t)do_lo;
        const uint64_t kd = dKm | (uint64_t)(k_lo + st * (ST >> 4));
        const uint64_t vd = dKm | (uint64_t)(v_lo + st * (ST >> 4));
#pragma unroll
        for (int k = 0; k < AB_D / 16; ++k) umma_ss(tmem_base, qd + 2 * k, kd + 2 * k, IDESC_S, k ? 1u : 0u);
#pragma unroll
        for (int k = 0; k < AB_D / 16; ++k) umma_ss(tmem_base + C_DP, od + 2 * k, vd + 2 * k, IDESC_S, k ? 1u : 0u);
        umma_commit(s_full);
        mbar_wait(p_full, j & 1);
        tc_fence_after();
        const uint64_t kmn = dMN | (uint64_t)(k_lo + st * (ST >> 4));
#pragma unroll
        for (int k = 0; k < KB / 16; ++k)
          umma_ts(tmem_base + C_DQ, tmem_base + 8 * k, kmn + 128 * k, IDESC_O, (j | k) != 0 ? 1u : 0u);
        umma_commit(&kv_empty[st]);
      }
      umma_commit(o_done);
    }
  } else {
    const int row = warp * 32 + lane;
    const int r = q0 + row;
    const uint32_t lane_off = static_cast<uint32_t>(warp * 32) << 16;
    const uint32_t tS = tmem_base + lane_off, tP = tmem_base + C_DP + lane_off, tO = tmem_base + C_DQ + lane_off;
    const long long sidx = ((long long)b * heads + h) * rows + r;
    const float my_lse = r < rows ? lse2[sidx] : INFINITY;  // rows past the end: P = exp2(-inf) = 0
    const float my_D = r < rows ? Dr[sidx] : 0.0f;
    for (int j = 0; j < nkv; ++j) {
      mbar_wait(s_full, j & 1);
      tc_fence_after();
      const int valid = L - j * KB;
#pragma unroll
      for (int c0 = 0; c0 < KB; c0 += 32) {  // 32 columns at a time keeps the kernel at two CTAs' worth of registers
        float s[32], dp[32];
        tmem_ld_32x32(tS + c0, s);
        tmem_ld_32x32(tP + c0, dp);
        tmem_ld_wait();
#pragma unroll
        for (int i = 0; i < 32; ++i) {
          const float p = ab_exp2(fmaf(s[i], scale_log2, -my_lse));
          s[i] = (c0 + i < valid) ? p * (dp[i] - my_D) : 0.0f;  // padded keys carry nothing
        }
        uint32_t* pk = reinterpret_cast<uint32_t*>(s);
#pragma unroll
        for (int i = 0; i < 16; ++i) pk[i] = pack_bf16x2(s[2 * i], s[2 * i + 1]);
        tmem_st_32x8(tS + c0 / 2, s);      // dS as bf16 pairs over columns [c0 / 2, c0 / 2 + 16): already read above
        tmem_st_32x8(tS + c0 / 2 + 8, s + 8);
      }
      tmem_st_wait();
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(p_full);
    }
    float* op = dqkv + ((long long)b * rows + r) * 3 * inner + h * AB_D;
    if (nkv > 0) {
      mbar_wait(o_done, 0);
      tc_fence_after();
#pragma unroll
      for (int c = 0; c < AB_D; c += 32) {
        float o[32];
        tmem_ld_32x32(tO + c, o);
        tmem_ld_wait();
        if (r < rows) {
#pragma unroll
          for (int d = 0; d < 32; d += 4)
            *reinterpret_cast<float4*>(op + c + d) =
                make_float4(o[d] * scale, o[d + 1] * scale, o[d + 2] * scale, o[d + 3] * scale);
        }
      }
    } else if (r < rows) {
      for (int c = 0; c < AB_D; c += 4) *reinterpret_cast<float4*>(op + c) = make_float4(0.f, 0.f, 0.f, 0.f);
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 5) {
    tc_fence_after();
    tmem_dealloc(tmem_base, AbCfg<KB>::TMEM_COLS);
  }
}

// ------------------------------------------------------------------------------------------------------------------
// dK, dV: grid (key tiles, heads, batches).  TMEM: S^T [0,KB) (P^T bf16 pairs over its first half), dP^T [KB,2KB)
// (dS^T bf16 pairs over its first half), dV [2KB,2KB+64), dK [2KB+64,2KB+128).  Lane = key, column = query of the step.
// ------------------------------------------------------------------------------------------------------------------
template <int KB>
__global__ void __launch_bounds__(192, KB == 64 ? 2 : 1)
tc_attn_dkv_kernel(const __grid_constant__ CUtensorMap tm_qkv, const __grid_constant__ CUtensorMap tm_q,
                   const __grid_constant__ CUtensorMap tm_do,
                   const float* __restrict__ lse2, const float* __restrict__ Dr, float* __restrict__ dqkv, int rows,
                   int heads, const int* __restrict__ kv_len, float scale, float scale_log2) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
  uint8_t* sK = smem;
  uint8_t* sV = sK + AB_TILE;
  constexpr int ST = AbCfg<KB>::STREAM_TILE;
  // TMEM columns: S^T [0, KB), dP^T [KB, 2 KB), dV [2 KB, 2 KB + 64), dK [2 KB + 64, 2 KB + 128)
  constexpr uint32_t C_DP = KB, C_DV = 2 * KB, C_DK = 2 * KB + 64;
  uint8_t* sQ = sV + AB_TILE;         // AB_STAGES
  uint8_t* sDO = sQ + AB_STAGES * ST;  // AB_STAGES
  uint64_t* bars = reinterpret_cast<uint64_t*>(sDO + AB_STAGES * ST);
  uint64_t* kv_full = bars;                // [1]
  uint64_t* q_full = bars + 1;             // [AB_STAGES]
  uint64_t* q_empty = q_full + AB_STAGES;  // [AB_STAGES]
  uint64_t* s_full = q_empty + AB_STAGES;
  uint64_t* p_full = s_full + 1;
  uint64_t* o_done = p_full + 1;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(o_done + 1);
  float* s_lse = reinterpret_cast<float*>(bars + 32);  // [2][KB]
  float* s_D = s_lse + 2 * KB;                         // [2][KB]

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int k0 = blockIdx.x * 128, h = blockIdx.y, b = blockIdx.z;
  const int inner = heads * AB_D;
  const int L = kv_len ? kv_len[b] : rows;
  const int nq = k0 < L ? (rows + KB - 1) / KB : 0;  // a tile of padded keys receives no gradient

  if (threadIdx.x == 0) {
    prefetch_tmap(&tm_qkv);
    prefetch_tmap(&tm_do);
    prefetch_tmap(&tm_q);
    mbar_init(kv_full, 1);
    for (int i = 0; i < AB_STAGES; ++i) {
      mbar_init(&q_full[i], 1);
      mbar_init(&q_empty[i], 1);
    }
    mbar_init(s_full, 1);
    mbar_init(p_full, 4);
    mbar_init(o_done, 1);
    fence_barrier_init();
  }
  if (warp == 5) tmem_alloc(tmem_slot, AbCfg<KB>::TMEM_COLS);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 4) {
    if (elect_one() && nq > 0) {
      mbar_expect_tx(kv_full, 2 * AB_TILE);
      tma_load_3d(sK, &tm_qkv, kv_full, inner + h * AB_D, k0, b);
      tma_load_3d(sV, &tm_qkv, kv_full, 2 * inner + h * AB_D, k0, b);
      for (int i = 0; i < nq; ++i) {
        const int st = i % AB_STAGES;
        mbar_wait(&q_empty[st], ((i / AB_STAGES) & 1) ^ 1);
        mbar_expect_tx(&q_full[st], 2 * ST);
        tma_load_3d(sQ + st * ST, &tm_q, &q_full[st], h * AB_D, i * KB, b);
        tma_load_3d(sDO + st * ST, &tm_do, &q_full[st], h * AB_D, i * KB, b);
      }
    }
  } else if (warp == 5) {
    if (elect_one() && nq > 0) {
      constexpr uint32_t IDESC_S = umma_idesc_bf16(128, KB, 0, 0);
      constexpr uint32_t IDESC_O = umma_idesc_bf16(128, 64, 0, 1);
      const uint64_t dKm = umma_desc_sw128(0, 16, 1024);
      const uint64_t dMN = umma_desc_sw128(0, 1024, 1024);
      const uint32_t k_lo = smem_u32(sK) >> 4, v_lo = smem_u32(sV) >> 4, q_lo = smem_u32(sQ) >> 4,
                     do_lo = smem_u32(sDO) >> 4;
      mbar_wait(kv_full, 0);
      for (int i = 0; i < nq; ++i) {
        const int st = i % AB_STAGES;
        mbar_wait(&q_full[st], (i / AB_STAGES) & 1);
        tc_fence_after();
        const uint64_t kd = dKm | (uint64_t)k_lo, vd = dKm | (uint64_t)v_lo;
        const uint64_t qd = dKm | (uint64_t)(q_lo + st * (ST >> 4));
        const uint64_t od = dKm | (uint64_t)(do_lo + st * (ST >> 4));
#pragma unroll
        for (int k = 0; k < AB_D / 16; ++k) umma_ss(tmem_base, kd + 2 * k, qd + 2 * k, IDESC_S, k ? 1u : 0u);       // S^T
#pragma unroll
        for (int k = 0; k < AB_D / 16; ++k) umma_ss(tmem_base + C_DP, vd + 2 * k, od + 2 * k, IDESC_S, k ? 1u : 0u);  // dP^T
        umma_commit(s_full);
        mbar_wait(p_full, i & 1);
        tc_fence_after();
        const uint64_t omn = dMN | (uint64_t)(do_lo + st * (ST >> 4));
        const uint64_t qmn = dMN | (uint64_t)(q_lo + st * (ST >> 4));
#pragma unroll
        for (int k = 0; k < KB / 16; ++k)  // dV += P^T dO_i
          umma_ts(tmem_base + C_DV, tmem_base + 8 * k, omn + 128 * k, IDESC_O, (i | k) != 0 ? 1u : 0u);
#pragma unroll
        for (int k = 0; k < KB / 16; ++k)  // dK += dS^T Q_i
          umma_ts(tmem_base + C_DK, tmem_base + C_DP + 8 * k, qmn + 128 * k, IDESC_O, (i | k) != 0 ? 1u : 0u);
        umma_commit(&q_empty[st]);
      }
      umma_commit(o_done);
    }
  } else {
    const int row = warp * 32 + lane;  // key inside the tile = TMEM lane
    const int kr = k0 + row;
    const bool key_ok = kr < L;
    const uint32_t lane_off = static_cast<uint32_t>(warp * 32) << 16;
    const uint32_t tS = tmem_base + lane_off, tP = tmem_base + C_DP + lane_off;
    const long long sbase = ((long long)b * heads + h) * rows;
    // per-query statistics of the step's KB queries: loaded from global memory ONE STEP AHEAD into registers (the
    // profile showed the load -> shared store -> barrier of the current step as the largest stall of the kernel), then
    // published through shared memory (double-buffered; the named barrier orders writers and readers)
    float nxt_lse = INFINITY, nxt_D = 0.0f;
    if (row < KB && nq > 0 && row < rows) {
      nxt_lse = lse2[sbase + row];
      nxt_D = Dr[sbase + row];
    }
    for (int i = 0; i < nq; ++i) {
      {
        if (row < KB) {
          s_lse[(i & 1) * KB + row] = nxt_lse;
          s_D[(i & 1) * KB + row] = nxt_D;
        }
        named_bar_sync(1, 128);
        if (row < KB && i + 1 < nq) {
          const int qr = (i + 1) * KB + row;
          nxt_lse = qr < rows ? lse2[sbase + qr] : INFINITY;
          nxt_D = qr < rows ? Dr[sbase + qr] : 0.0f;
        }
      }
      const float4* l4 = reinterpret_cast<const float4*>(s_lse + (i & 1) * KB);
      const float4* d4 = reinterpret_cast<const float4*>(s_D + (i & 1) * KB);
      mbar_wait(s_full, i & 1);
      tc_fence_after();
#pragma unroll
      for (int c0 = 0; c0 < KB; c0 += 32) {
        float s[32], dp[32];
        tmem_ld_32x32(tS + c0, s);
        tmem_ld_32x32(tP + c0, dp);
        tmem_ld_wait();
#pragma unroll
        for (int c = 0; c < 32; c += 4) {
          const float4 lv = l4[(c0 + c) >> 2], dv = d4[(c0 + c) >> 2];
          const float ls[4] = {lv.x, lv.y, lv.z, lv.w}, ds_[4] = {dv.x, dv.y, dv.z, dv.w};
#pragma unroll
          for (int q = 0; q < 4; ++q) {
            const float p = key_ok ? ab_exp2(fmaf(s[c + q], scale_log2, -ls[q])) : 0.0f;
            s[c + q] = p;
            dp[c + q] = key_ok ? p * (dp[c + q] - ds_[q]) : 0.0f;
          }
        }
        uint32_t* pk = reinterpret_cast<uint32_t*>(s);
        uint32_t* dk = reinterpret_cast<uint32_t*>(dp);
#pragma unroll
        for (int c = 0; c < 16; ++c) {
          pk[c] = pack_bf16x2(s[2 * c], s[2 * c + 1]);
          dk[c] = pack_bf16x2(dp[2 * c], dp[2 * c + 1]);
        }
        tmem_st_32x8(tS + c0 / 2, s);  // P^T and dS^T as bf16 pairs over columns that have been read already
        tmem_st_32x8(tS + c0 / 2 + 8, s + 8);
        tmem_st_32x8(tP + c0 / 2, dp);
        tmem_st_32x8(tP + c0 / 2 + 8, dp + 8);
      }
      tmem_st_wait();
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(p_full);
    }
    float* opk = dqkv + ((long long)b * rows + kr) * 3 * inner + inner + h * AB_D;
    float* opv = opk + inner;
    if (nq > 0) {
      mbar_wait(o_done, 0);
      tc_fence_after();
#pragma unroll
      for (int c = 0; c < AB_D; c += 32) {
        float o[32];
        tmem_ld_32x32(tmem_base + C_DV + lane_off + c, o);
        tmem_ld_wait();
        if (kr < rows) {
#pragma unroll
          for (int d = 0; d < 32; d += 4) *reinterpret_cast<float4*>(opv + c + d) = make_float4(o[d], o[d + 1], o[d + 2], o[d + 3]);
        }
        tmem_ld_32x32(tmem_base + C_DK + lane_off + c, o);
        tmem_ld_wait();
        if (kr < rows) {
#pragma unroll
          for (int d = 0; d < 32; d += 4)
            *reinterpret_cast<float4*>(opk + c + d) =
                make_float4(o[d] * scale, o[d + 1] * scale, o[d + 2] * scale, o[d + 3] * scale);
        }
      }
    } else if (kr < rows) {
      for (int c = 0; c < AB_D; c += 4) {
        *reinterpret_cast<float4*>(opk + c) = make_float4(0.f, 0.f, 0.f, 0.f);
        *reinterpret_cast<float4*>(opv + c) = make_float4(0.f, 0.f, 0.f, 0.f);
      }
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 5) {
    tc_fence_after();
    tmem_dealloc(tmem_base, AbCfg<KB>::TMEM_COLS);
  }
}

// d_o16 = bf16(d_o) and D[b,h,r] = sum_d d_o o over the head's 64 columns: 16 lanes per (row, head), float4 each.
__global__ void __launch_bounds__(256)
attn_bwd_prep_kernel(const float* __restrict__ o, const float* __restrict__ d_o, int rows, int heads, long long total16,
                     bf16* __restrict__ d_o16, float* __restrict__ D) {
  const long long g = (long long)blockIdx.x * 256 + threadIdx.x;  // one float4 of one (row, head)
  const bool live = g < total16;
  float acc = 0.0f;
  if (live) {
    const float4 a = reinterpret_cast<const float4*>(o)[g];
    const float4 d = reinterpret_cast<const float4*>(d_o)[g];
    acc = a.x * d.x + a.y * d.y + a.z * d.z + a.w * d.w;
    uint2 pk;
    pk.x = pack_bf16x2(d.x, d.y);
    pk.y = pack_bf16x2(d.z, d.w);
    reinterpret_cast<uint2*>(d_o16)[g] = pk;
  }
#pragma unroll
  for (int off = 8; off > 0; off >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, off, 16);
  if (live && (threadIdx.x & 15) == 0) {
    const long long rh = g >> 4;  // (global row) * heads + head
    const long long m = rh / heads;
    const int h = (int)(rh - m * heads);
    const long long bb = m / rows;
    D[(bb * heads + h) * rows + (m - bb * rows)] = acc;
  }
}

void attn_bwd_prep(cudaStream_t st, const float* o, const float* d_o, int batches, int rows, int heads, bf16* d_o16,
                   float* D) {
  const long long total16 = (long long)batches * rows * heads * 16;
  attn_bwd_prep_kernel<<<(unsigned)((total16 + 255) / 256), 256, 0, st>>>(o, d_o, rows, heads, total16, d_o16, D);
  F5B_POST_LAUNCH();
}

template <int KB>
static void launch_attn_bwd(cudaStream_t st, const bf16* qkv, const bf16* d_o, const float* lse2, const float* D,
                            int batches, int rows, int heads, const int* kv_len, float scale, float* dqkv) {
  const int inner = heads * AB_D;
  CUtensorMap tq, tqs, td, tds;  // resident tiles: 128-row boxes; streamed tiles: KB-row boxes
  {
    uint64_t dims[3] = {(uint64_t)3 * inner, (uint64_t)rows, (uint64_t)batches};
    uint64_t strides[2] = {(uint64_t)3 * inner * 2, (uint64_t)3 * inner * 2 * (uint64_t)rows};
    uint32_t box[3] = {AB_D, 128, 1}, boxs[3] = {AB_D, KB, 1};
    make_tmap_bf16(&tq, qkv, 3, dims, strides, box);
    make_tmap_bf16(&tqs, qkv, 3, dims, strides, boxs);
  }
  {
    uint64_t dims[3] = {(uint64_t)inner, (uint64_t)rows, (uint64_t)batches};
    uint64_t strides[2] = {(uint64_t)inner * 2, (uint64_t)inner * 2 * (uint64_t)rows};
    uint32_t box[3] = {AB_D, 128, 1}, boxs[3] = {AB_D, KB, 1};
    make_tmap_bf16(&td, d_o, 3, dims, strides, box);
    make_tmap_bf16(&tds, d_o, 3, dims, strides, boxs);
  }
  dim3 grid((rows + 127) / 128, heads, batches);
  const float sl2 = scale * 1.4426950408889634f;
  tc_attn_dq_kernel<KB><<<grid, 192, AbCfg<KB>::SMEM, st>>>(tq, tqs, td, lse2, D, dqkv, rows, heads, kv_len, scale, sl2);
  F5B_POST_LAUNCH();
  tc_attn_dkv_kernel<KB><<<grid, 192, AbCfg<KB>::SMEM, st>>>(tq, tqs, tds, lse2, D, dqkv, rows, heads, kv_len, scale, sl2);
  F5B_POST_LAUNCH();
}

// F5B_ATTN_BWD_KB = 64 (default: two CTAs per SM) | 128 (one CTA per SM; kept for comparison)
void tc_attention_backward(cudaStream_t st, const bf16* qkv, const bf16* d_o, const float* lse2, const float* D,
                           int batches, int rows, int heads, const int* kv_len, float scale, float* dqkv) {
  static const int kb = [] {
    const char* e = std::getenv("F5B_ATTN_BWD_KB");
    return e ? std::atoi(e) : 64;
  }();
  if (kb == 128) launch_attn_bwd<128>(st, qkv, d_o, lse2, D, batches, rows, heads, kv_len, scale, dqkv);
  else launch_attn_bwd<64>(st, qkv, d_o, lse2, D, batches, rows, heads, kv_len, scale, dqkv);
}

void tc_attn_bwd_init_kernels() {
  F5B_CHECK_CUDA(cudaFuncSetAttribute(tc_attn_dq_kernel<64>, cudaFuncAttributeMaxDynamicSharedMemorySize, AbCfg<64>::SMEM));
  F5B_CHECK_CUDA(cudaFuncSetAttribute(tc_attn_dkv_kernel<64>, cudaFuncAttributeMaxDynamicSharedMemorySize, AbCfg<64>::SMEM));
  F5B_CHECK_CUDA(cudaFuncSetAttribute(tc_attn_dq_kernel<128>, cudaFuncAttributeMaxDynamicSharedMemorySize, AbCfg<128>::SMEM));
  F5B_CHECK_CUDA(cudaFuncSetAttribute(tc_attn_dkv_kernel<128>, cudaFuncAttributeMaxDynamicSharedMemorySize, AbCfg<128>::SMEM));
}

}  // namespace f5b
