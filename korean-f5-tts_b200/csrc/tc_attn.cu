// Non-causal multi-head attention, head_dim 64, on tcgen05 tensor cores with TMEM accumulators.
//
// Replaces AttnProcessor.__call__'s F.scaled_dot_product_attention (reference src/f5_tts/model/modules.py:499-508;
// flash-attn variant :510-531) for the DiT / UNetT blocks.  q, k (already rotated by the QKV GEMM epilogue) and v
// live in one bf16 [batches*rows, 3*heads*64] buffer.  One CTA (one per SM) owns TWO 128-query tiles of one
// (head, batch element) and streams the keys / values once for both:
//
//   warps 0..3 : softmax of query tile 0 -- thread t owns query row t (= TMEM lane t)
//   warps 4..7 : softmax of query tile 1
//   warp 8     : TMA producer (Q0, Q1 once; K_j, V_j through a 3-stage ring)
//   warp 9     : MMA issuer.  Per key block j and tile t:  S_t = Q_t K_j^T (4 x 128x128x16) into TMEM,
//                O_t += P_t V_j (8 x 128x64x16, A = P_t read from TMEM, V read MN-major) accumulated IN TMEM.
//
// Each S block is read from TMEM exactly once (128 registers per thread) and P goes back to tensor memory as packed
// bf16 pairs (tcgen05.st), from where the P*V MMA takes it as its A operand.  O is never read back per block: the
// running maximum is only raised when a block exceeds it by more than 2^8 (lazy rescaling; P is then still < 256,
// harmless in bf16 / fp32), in which case the warp rescales its O rows in TMEM (tcgen05.ld / st).  The final O is read
// once, divided by the row sum and stored.
//
// Head dim 64 makes the softmax, not the tensor pipe, the floor: one exponential per 256 flop, 16 MUFU lanes per SM
// against 8192 flop per clock, i.e. 2048 MUFU cycles per 128-key block of a tile pair against 1024 tensor cycles.
// What the kernel does about it (round 2; each step is a template switch so it can be measured against the previous):
//   SEQ   : the two tiles' exponential sections are SEQUENCED per warp pair (warps w and w+4 share a scheduler and
//           its MUFU): without it both tiles drift into lock step -- both exponentiate at half rate, then both wait
//           for the tensor core -- which the in-kernel timeline showed as a 3000-cycle period; with the hand-over one
//           tile's exponentials run while the other tile does everything else (TMEM load, row maximum, pack, store,
//           S / P*V MMAs).
//   EARLY : P lives in its own TMEM columns instead of over S, so S_t(j+1) is issued as soon as S_t(j) has been read
//           into registers and its MMA time leaves the tile's dependency chain.
//   POLY  : POLY of every 8 elements take exp2 on the FMA pipe instead of the MUFU (Cody-Waite split by a
//           magic-number add, degree-3 minimax polynomial on [-0.5, 0.5], max relative error 7.6e-5 -- 50x below the
//           bf16 rounding of P -- exponent inserted with one LEA).
//   All softmax arithmetic is scalar: packed f32x2 forms issue at half the per-element rate (tools/ubench/pipes.cu).
//   Measured (profiles/r2_attn_variants.log): SEQ wins 5-7 %, EARLY and POLY lose; default variant = 100.
#include "common.cuh"
#include "sm100_ptx.cuh"
#include "tc_gemm.h"

#include <cstdlib>

namespace f5b {

using namespace ptx;

static constexpr int AT_BQ = 128, AT_BKV = 128, AT_D = 64;
static constexpr int AT_TILE_BYTES = 128 * 64 * 2;  // 16 KB: Q_t, K_j, V_j
static constexpr int AT_KV_STAGES = 3;
static constexpr int AT_SMEM = AT_TILE_BYTES * (2 + 2 * AT_KV_STAGES) + 1024 + 512;
// TMEM columns: S0 [0,128) S1 [128,256) O0 [256,320) O1 [320,384) and, with EARLY, P0 [384,448) P1 [448,512)
static constexpr uint32_t AT_TMEM_COLS = 512;
static constexpr unsigned AT_STAGGER_NS = 700;  // legacy variant only: ~half of its per-block period
static constexpr float AT_LAZY = 8.0f;          // rescale only when the block maximum exceeds the reference by 2^8

// Debug timeline (f5b_debug_attn_trace): CTA (0,0,0) stamps clock64 at its pipeline events.  [who][j][event]
__device__ long long g_attn_trace[3 * 64 * 8];
__device__ int g_attn_trace_on = 0;
#define AT_STAMP(who, j, ev)                                                                         \
  do {                                                                                               \
    if (trace && (j) < 64) g_attn_trace[((who) * 64 + (j)) * 8 + (ev)] = clock64();                  \
  } while (0)

__device__ __forceinline__ float fast_exp2(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}


// 2^x on the FMA pipe.  n = round(x) sits in the low mantissa bits of x + 1.5 * 2^23; f = x - n lies in [-0.5, 0.5];
// 2^f by a degree-3 minimax polynomial (relative error <= 7.6e-5, 50x below the bf16 rounding of P); the exponent is
// inserted by adding n << 23 to the bits (the magic constant's own bits shift out).  x is clamped at -126 (padded keys
// are -inf).  Scalar on purpose: measured on B200 (tools/ubench/pipes.cu) FFMA2 / FADD2 issue at ~4 cycles per warp
// instruction, i.e. HALF the per-element rate of scalar FFMA at ~1 cycle -- the packed forms only save issue slots.
__device__ __forceinline__ float exp2_poly(float x) {
  x = fmaxf(x, -126.0f);
  const float xr = x + 12582912.0f;
  const float f = x - (xr - 12582912.0f);
  float p = fmaf(f, 0.05520550534129143f, 0.24261397123336792f);
  p = fmaf(p, f, 0.6932547688484192f);
  p = fmaf(p, f, 0.9999276995658875f);
  return __int_as_float(__float_as_int(p) + (__float_as_int(xr) << 23));
}

template <int SEQ, int EARLY, int POLY>
__global__ void __launch_bounds__(320, 1)
tc_attn_kernel(const __grid_constant__ CUtensorMap tm, bf16* __restrict__ out, int rows, int heads,
               const int* __restrict__ kv_len, float scale_log2, int out_f16, float* __restrict__ lse2_out) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
  uint8_t* sQ = smem;                                  // 2 tiles
  uint8_t* sK = sQ + 2 * AT_TILE_BYTES;                // AT_KV_STAGES
  uint8_t* sV = sK + AT_KV_STAGES * AT_TILE_BYTES;     // AT_KV_STAGES
  uint64_t* bars = reinterpret_cast<uint64_t*>(sV + AT_KV_STAGES * AT_TILE_BYTES);
  uint64_t* q_full = bars;                        // [1]
  uint64_t* kv_full = bars + 1;                   // [AT_KV_STAGES]
  uint64_t* kv_empty = kv_full + AT_KV_STAGES;    // [AT_KV_STAGES]
  uint64_t* s_full = kv_empty + AT_KV_STAGES;     // [2]  MMA -> softmax: S_t(j) is in TMEM
  uint64_t* p_full = s_full + 2;                  // [2]  softmax -> MMA: P_t(j) is in TMEM
  uint64_t* o_done = p_full + 2;                  // [2]  MMA -> softmax: P_t V (of the last block, or of every block with EARLY) has completed
  uint64_t* s_free = o_done + 2;                  // [2]  EARLY: softmax -> MMA: S_t(j) is in registers, the columns may be overwritten
  uint64_t* seq = s_free + 2;                     // [2][4] SEQ: warp w of tile t may start its exponentials
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(seq + 8);

  KT_DECL
  KT_BEGIN();
  pdl_trigger();
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int q0 = blockIdx.x * 2 * AT_BQ, h = blockIdx.y, b = blockIdx.z;
  const int inner = heads * AT_D;
  const int L = kv_len ? kv_len[b] : rows;
  const int nkv = (L + AT_BKV - 1) / AT_BKV;
  const bool trace = g_attn_trace_on && blockIdx.x == 0 && blockIdx.y == 0 && blockIdx.z == 0 && lane == 0 &&
                     (warp == 0 || warp == 4 || warp == 9);

  if (threadIdx.x == 0) {
    prefetch_tmap(&tm);
    mbar_init(q_full, 1);
    for (int i = 0; i < AT_KV_STAGES; ++i) {
      mbar_init(&kv_full[i], 1);
      mbar_init(&kv_empty[i], 1);
    }
    for (int t = 0; t < 2; ++t) {
      mbar_init(&s_full[t], 1);
      mbar_init(&p_full[t], 4);
      mbar_init(&o_done[t], 1);
      mbar_init(&s_free[t], 4);
    }
    for (int i = 0; i < 8; ++i) mbar_init(&seq[i], 1);
    fence_barrier_init();
  }
  if (warp == 9) tmem_alloc(tmem_slot, AT_TMEM_COLS);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  pdl_wait();  // qkv comes from the previous kernel; everything above overlapped its tail
  KT_WAITED();

  if (warp == 8) {
    if (elect_one()) {
      // ------------------------------- TMA producer -------------------------------
      mbar_expect_tx(q_full, 2 * AT_TILE_BYTES);
      tma_load_3d(sQ, &tm, q_full, h * AT_D, q0, b);
      tma_load_3d(sQ + AT_TILE_BYTES, &tm, q_full, h * AT_D, q0 + AT_BQ, b);
      for (int j = 0; j < nkv; ++j) {
        const int st = j % AT_KV_STAGES;
        mbar_wait(&kv_empty[st], ((j / AT_KV_STAGES) & 1) ^ 1);
        mbar_expect_tx(&kv_full[st], 2 * AT_TILE_BYTES);
        tma_load_3d(sK + st * AT_TILE_BYTES, &tm, &kv_full[st], inner + h * AT_D, j * AT_BKV, b);
        tma_load_3d(sV + st * AT_TILE_BYTES, &tm, &kv_full[st], 2 * inner + h * AT_D, j * AT_BKV, b);
      }
    }
  } else if (warp == 9) {
    // elect.sync, not `lane == 0`: the compiler then KNOWS a single lane is active and issues each tcgen05.mma straight
    // from uniform registers; behind a lane test it wraps every MMA in a loop over the possibly-active lanes
    // (ELECT / R2UR / BRA.U.ANY, ~70 cycles per MMA -- twice the 32-cycle floor of a 128x64x16 P*V instruction)
    if (elect_one()) {
      // ------------------------------- MMA issuer ---------------------------------
      constexpr uint32_t IDESC_S = umma_idesc_bf16(128, 128, 0, 0);
      constexpr uint32_t IDESC_O = umma_idesc_bf16(128, 64, 0, 1);  // B = V is MN-major
      // descriptors: only the 14-bit start-address field changes; built once, advanced by adds (single issuing lane)
      const uint64_t dK = umma_desc_sw128(0, 16, 1024);    // K-major operands (Q, K_j)
      const uint64_t dV = umma_desc_sw128(0, 1024, 1024);  // MN-major operand (V_j)
      const uint32_t q_lo = smem_u32(sQ) >> 4, k_lo = smem_u32(sK) >> 4, v_lo = smem_u32(sV) >> 4;
      const uint32_t p_col = EARLY ? 384u : 0u;            // P_t: own columns, or over the first half of S_t
      const uint32_t p_stride = EARLY ? 64u : 128u;
      auto issue_s = [&](int t, int st) {
        const uint64_t qd = dK | (uint64_t)(q_lo + t * (AT_TILE_BYTES >> 4));
        const uint64_t kd = dK | (uint64_t)(k_lo + st * (AT_TILE_BYTES >> 4));
        umma_ss(tmem_base + t * 128, qd, kd, IDESC_S, 0u);
#pragma unroll
        for (int k = 1; k < AT_D / 16; ++k) umma_ss(tmem_base + t * 128, qd + 2 * k, kd + 2 * k, IDESC_S, 1u);
        umma_commit(&s_full[t]);
      };
      auto issue_pv = [&](int t, int st, int j) {
        const uint64_t vd = dV | (uint64_t)(v_lo + st * (AT_TILE_BYTES >> 4));
#pragma unroll
        for (int k = 0; k < AT_BKV / 16; ++k)  // A = P_t in TMEM: 8 columns (16 keys) per k-step; V: 16 rows
          umma_ts(tmem_base + 256 + t * 64, tmem_base + p_col + t * p_stride + 8 * k, vd + 128 * k, IDESC_O,
                  (j | k) != 0 ? 1u : 0u);
      };
      mbar_wait(q_full, 0);
      mbar_wait(&kv_full[0], 0);
      tc_fence_after();
      issue_s(0, 0);
      issue_s(1, 0);
      for (int j = 0; j < nkv; ++j) {
        const int st = j % AT_KV_STAGES;
        const int stn = (j + 1) % AT_KV_STAGES;
        if (j + 1 < nkv) mbar_wait(&kv_full[stn], ((j + 1) / AT_KV_STAGES) & 1);
        if (EARLY) {
          // S_t(j+1) does not wait for P_t(j): it only needs S_t(j) to have left its columns
          if (j + 1 < nkv) {
            for (int t = 0; t < 2; ++t) {
              mbar_wait(&s_free[t], j & 1);
              tc_fence_after();
              issue_s(t, stn);
            }
          }
          for (int t = 0; t < 2; ++t) {
            AT_STAMP(2, j, t * 3);
            mbar_wait(&p_full[t], j & 1);
            tc_fence_after();
            AT_STAMP(2, j, t * 3 + 1);
            issue_pv(t, st, j);
            umma_commit(&o_done[t]);  // every block: the tile's O may be rescaled, its P columns rewritten
            AT_STAMP(2, j, t * 3 + 2);
          }
        } else {
          for (int t = 0; t < 2; ++t) {
            AT_STAMP(2, j, t * 3);
            mbar_wait(&p_full[t], j & 1);
            tc_fence_after();
            AT_STAMP(2, j, t * 3 + 1);
            issue_pv(t, st, j);
            if (j + 1 < nkv) issue_s(t, stn);  // S_t(j+1): its commit also covers PV_t(j)
            else umma_commit(&o_done[t]);
            AT_STAMP(2, j, t * 3 + 2);
          }
        }
        umma_commit(&kv_empty[st]);  // K_j / V_j have been read by every MMA issued so far
        AT_STAMP(2, j, 6);
      }
    }
  } else {
    // ------------------------------- softmax / output ---------------------------
    const int t = warp >> 2;                 // query tile of this warp group
    const int wq = warp & 3;                 // scheduler (and MUFU) shared with warp wq of the other tile
    const int row = wq * 32 + lane;          // row inside the tile = TMEM lane
    const uint32_t lane_off = static_cast<uint32_t>(wq * 32) << 16;
    const uint32_t tS = tmem_base + t * 128 + lane_off;
    const uint32_t tO = tmem_base + 256 + t * 64 + lane_off;
    const uint32_t tP = EARLY ? tmem_base + 384 + t * 64 + lane_off : tS;
    float m_ref = -INFINITY;
    float l = 0.0f;
    if (!SEQ && t == 1 && nkv > 1) __nanosleep(AT_STAGGER_NS);  // legacy: start tile 1 half a period late

    for (int j = 0; j < nkv; ++j) {
      AT_STAMP(t, j, 0);
      mbar_wait(&s_full[t], j & 1);
      tc_fence_after();
      AT_STAMP(t, j, 1);
      float s[AT_BKV];
      tmem_ld_32x32(tS, s);
      tmem_ld_32x32(tS + 32, s + 32);
      tmem_ld_32x32(tS + 64, s + 64);
      tmem_ld_32x32(tS + 96, s + 96);
      tmem_ld_wait();
      AT_STAMP(t, j, 2);
      if (EARLY) {
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(&s_free[t]);
      }
      const int valid = L - j * AT_BKV;  // key columns >= valid are padding (only in the last block)
      if (valid < AT_BKV) {
#pragma unroll
        for (int i = 0; i < AT_BKV; ++i)
          if (i >= valid) s[i] = -INFINITY;
      }
      float bm4[4] = {s[0], s[1], s[2], s[3]};  // four independent chains: the reduction is latency-, not issue-bound
#pragma unroll
      for (int i = 4; i < AT_BKV; i += 4) {
        bm4[0] = fmaxf(bm4[0], s[i]);
        bm4[1] = fmaxf(bm4[1], s[i + 1]);
        bm4[2] = fmaxf(bm4[2], s[i + 2]);
        bm4[3] = fmaxf(bm4[3], s[i + 3]);
      }
      const float bm = fmaxf(fmaxf(bm4[0], bm4[1]), fmaxf(bm4[2], bm4[3]));
      if (EARLY && j > 0) {  // P_t V(j-1) done: O_t may be rescaled and the P columns rewritten (long since true)
        mbar_wait(&o_done[t], (j - 1) & 1);
        tc_fence_after();
      }
      if (j == 0) {
        m_ref = bm;
      } else {
        const bool need = (bm - m_ref) * scale_log2 > AT_LAZY;
        if (__any_sync(0xffffffffu, need)) {
          // raise the reference maximum of the rows that need it and rescale their O accumulators in TMEM
          // (PV_t(j-1) is complete: the commit that raised s_full covers it, or o_done was just observed)
          const float alpha = need ? fast_exp2((m_ref - bm) * scale_log2) : 1.0f;
          if (need) m_ref = bm;
          l *= alpha;
#pragma unroll 1
          for (int c = 0; c < AT_D; c += 8) {
            float o[8];
            tmem_ld_32x8(tO + c, o);
            tmem_ld_wait();
#pragma unroll
            for (int i = 0; i < 8; ++i) o[i] *= alpha;
            tmem_st_32x8(tO + c, o);
          }
          tmem_st_wait();
        }
      }
      AT_STAMP(t, j, 3);
      if (SEQ && (t == 1 || j > 0)) mbar_wait(&seq[t * 4 + wq], (t == 1 ? j : j - 1) & 1);  // my turn on the MUFU
      AT_STAMP(t, j, 4);
      // Scalar arithmetic on purpose: one FFMA, one MUFU.EX2, one FADD per element overlap at ~8.6 cycles per element
      // and scheduler (MUFU-bound, tools/ubench/pipes.cu); POLY of every 8 elements take the FMA-pipe polynomial.
      const float moff = m_ref * scale_log2;
      float ps4[4] = {0.0f, 0.0f, 0.0f, 0.0f};
#pragma unroll
      for (int i = 0; i < AT_BKV; i += 4) {
#pragma unroll
        for (int q = 0; q < 4; ++q) {
          const float x = fmaf(s[i + q], scale_log2, -moff);
          s[i + q] = (((i + q) & 7) < POLY) ? exp2_poly(x) : fast_exp2(x);  // exp2(-inf) = 0 for padded keys
          ps4[q] += s[i + q];
        }
      }
      l += (ps4[0] + ps4[1]) + (ps4[2] + ps4[3]);
      {
        uint32_t* pk = reinterpret_cast<uint32_t*>(s);
#pragma unroll
        for (int i = 0; i < AT_BKV / 2; ++i) pk[i] = pack_bf16x2(s[2 * i], s[2 * i + 1]);
        tmem_st_32x32(tP, s);
        tmem_st_32x32(tP + 32, s + 32);
        AT_STAMP(t, j, 5);
        if (SEQ) {  // every exponential of this block has been issued: hand the MUFU to the other tile's warp
          __syncwarp();
          if (lane == 0) mbar_arrive(&seq[(1 - t) * 4 + wq]);
        }
        tmem_st_wait();
      }
      AT_STAMP(t, j, 6);
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&p_full[t]);
      AT_STAMP(t, j, 7);
    }

    if (nkv > 0) mbar_wait(&o_done[t], EARLY ? (nkv - 1) & 1 : 0);  // (no keys at all: O is never produced, the rows
    tc_fence_after();                                               //  below store 0 * inf = NaN like softmax over an
    const int r = q0 + t * AT_BQ + row;                             //  empty set in the reference)
    const float inv = 1.0f / l;
    // training: the row's log-sum-exp in the log2 domain, P = exp2(s * scale_log2 - lse2) (read by tc_attn_bwd.cu)
    if (lse2_out != nullptr && r < rows) lse2_out[((long long)b * heads + h) * rows + r] = fmaf(m_ref, scale_log2, log2f(l));
    bf16* op = out + ((long long)b * rows + r) * inner + h * AT_D;
#pragma unroll
    for (int c = 0; c < AT_D; c += 32) {
      float o[32];
      tmem_ld_32x32(tO + c, o);
      tmem_ld_wait();
      if (r < rows) {
#pragma unroll
        for (int d = 0; d < 32; d += 8) {
          uint4 pk;
          if (out_f16) {
            pk.x = pack_f16x2(o[d] * inv, o[d + 1] * inv);
            pk.y = pack_f16x2(o[d + 2] * inv, o[d + 3] * inv);
            pk.z = pack_f16x2(o[d + 4] * inv, o[d + 5] * inv);
            pk.w = pack_f16x2(o[d + 6] * inv, o[d + 7] * inv);
          } else {
            pk.x = pack_bf16x2(o[d] * inv, o[d + 1] * inv);
            pk.y = pack_bf16x2(o[d + 2] * inv, o[d + 3] * inv);
            pk.z = pack_bf16x2(o[d + 4] * inv, o[d + 5] * inv);
            pk.w = pack_bf16x2(o[d + 6] * inv, o[d + 7] * inv);
          }
          *reinterpret_cast<uint4*>(op + c + d) = pk;
        }
      }
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 9) {
    tc_fence_after();
    tmem_dealloc(tmem_base, AT_TMEM_COLS);
  }
  KT_END(2);
}

KT_DEFINE_SETTER(ktrace_set_tc_attn)

// Variant code = SEQ * 100 + EARLY * 10 + POLY (F5B_ATTN_VARIANT overrides the default; read at every launch so one
// process can time them against each other).  0 is the round-1 kernel (nanosleep stagger, scalar arithmetic aside).
static constexpr int AT_DEFAULT_VARIANT = 100;  // measured best on B200 at the C2 / C3 / C4 shapes (profiles/r2_attn_variants.log)
typedef void (*AttnKernel)(const CUtensorMap, bf16*, int, int, const int*, float, int, float*);
static AttnKernel attn_kernel_for(int v) {
  switch (v) {
    case 0: return tc_attn_kernel<0, 0, 0>;
    case 10: return tc_attn_kernel<0, 1, 0>;
    case 11: return tc_attn_kernel<0, 1, 1>;
    case 100: return tc_attn_kernel<1, 0, 0>;
    case 110: return tc_attn_kernel<1, 1, 0>;
    case 111: return tc_attn_kernel<1, 1, 1>;
    case 112: return tc_attn_kernel<1, 1, 2>;
    case 113: return tc_attn_kernel<1, 1, 3>;
    default: return nullptr;
  }
}
static const int kAttnVariants[] = {0, 10, 11, 100, 110, 111, 112, 113};

void tc_attention(cudaStream_t st, const bf16* qkv, bf16* out, int batches, int rows, int heads, const int* kv_len,
                  float scale, int out_f16, float* lse2_out) {
  const int inner = heads * AT_D;
  CUtensorMap tm;
  uint64_t dims[3] = {(uint64_t)3 * inner, (uint64_t)rows, (uint64_t)batches};
  uint64_t strides[2] = {(uint64_t)3 * inner * 2, (uint64_t)3 * inner * 2 * (uint64_t)rows};
  uint32_t box[3] = {AT_D, 128, 1};
  make_tmap_bf16(&tm, qkv, 3, dims, strides, box);
  dim3 grid((rows + 2 * AT_BQ - 1) / (2 * AT_BQ), heads, batches);
  const char* env = std::getenv("F5B_ATTN_VARIANT");
  AttnKernel k = attn_kernel_for(env ? std::atoi(env) : AT_DEFAULT_VARIANT);
  F5B_REQUIRE(k != nullptr, "unknown F5B_ATTN_VARIANT");
  launch_pdl_id(2, k, grid, dim3(320), AT_SMEM, st, tm, out, rows, heads, kv_len, scale * 1.4426950408889634f, out_f16,
                lse2_out);
  F5B_POST_LAUNCH();
}

void attn_trace_enable(int on) { F5B_CHECK_CUDA(cudaMemcpyToSymbol(g_attn_trace_on, &on, sizeof(int))); }
void attn_trace_read(long long* host) {
  F5B_CHECK_CUDA(cudaMemcpyFromSymbol(host, g_attn_trace, sizeof(long long) * 3 * 64 * 8));
}

void tc_init_kernels() {
  for (int v : kAttnVariants)
    F5B_CHECK_CUDA(cudaFuncSetAttribute(attn_kernel_for(v), cudaFuncAttributeMaxDynamicSharedMemorySize, AT_SMEM));
}

}  // namespace f5b
