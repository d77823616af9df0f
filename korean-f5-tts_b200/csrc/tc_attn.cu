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
// The two tiles ping-pong: while tile 0's threads exponentiate block j, the tensor core runs tile 1's S / PV, so
// neither the MMA latency nor the softmax is exposed.  Each S block is read from TMEM exactly once (128 registers per
// thread) and P goes back into the same TMEM columns as packed bf16 pairs (tcgen05.st), from where the P*V MMA takes
// it as its A operand.  O is never read back per block: the running maximum is only raised when a block exceeds it by more than
// 2^8 (lazy rescaling; P is then still < 256, harmless in bf16 / fp32), in which case the warp rescales its O rows in
// TMEM (tcgen05.ld / st).  The final O is read once, divided by the row sum and stored as bf16.
#include "common.cuh"
#include "sm100_ptx.cuh"
#include "tc_gemm.h"

namespace f5b {

using namespace ptx;

static constexpr int AT_BQ = 128, AT_BKV = 128, AT_D = 64;
static constexpr int AT_TILE_BYTES = 128 * 64 * 2;  // 16 KB: Q_t, K_j, V_j
static constexpr int AT_KV_STAGES = 3;
static constexpr int AT_SMEM = AT_TILE_BYTES * (2 + 2 * AT_KV_STAGES) + 1024 + 256;
static constexpr uint32_t AT_TMEM_COLS = 512;  // S0 [0,128) S1 [128,256) O0 [256,320) O1 [320,384)
static constexpr unsigned AT_STAGGER_NS = 700;  // ~1300 cycles: half of the measured per-block period
static constexpr float AT_LAZY = 8.0f;         // rescale only when the block maximum exceeds the reference by 2^8

// Debug timeline (f5b_debug_attn_trace): CTA (0,0,0) stamps clock64 at its pipeline events.  [who][j][event]
__device__ long long g_attn_trace[3 * 64 * 4];
__device__ int g_attn_trace_on = 0;
#define AT_STAMP(who, j, ev)                                                                         \
  do {                                                                                               \
    if (trace && (j) < 64) g_attn_trace[((who) * 64 + (j)) * 4 + (ev)] = clock64();                  \
  } while (0)

__device__ __forceinline__ float fast_exp2(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}

__global__ void __launch_bounds__(320, 1)
tc_attn_kernel(const __grid_constant__ CUtensorMap tm, bf16* __restrict__ out, int rows, int heads,
               const int* __restrict__ kv_len, float scale_log2, int out_f16) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
  uint8_t* sQ = smem;                                  // 2 tiles
  uint8_t* sK = sQ + 2 * AT_TILE_BYTES;                // AT_KV_STAGES
  uint8_t* sV = sK + AT_KV_STAGES * AT_TILE_BYTES;     // AT_KV_STAGES
  uint64_t* bars = reinterpret_cast<uint64_t*>(sV + AT_KV_STAGES * AT_TILE_BYTES);
  uint64_t* q_full = bars;                        // [1]
  uint64_t* kv_full = bars + 1;                   // [AT_KV_STAGES]
  uint64_t* kv_empty = kv_full + AT_KV_STAGES;    // [AT_KV_STAGES]
  uint64_t* s_full = kv_empty + AT_KV_STAGES;     // [2]  MMA -> softmax: S_t(j) is in TMEM (and PV_t(j-1) is done)
  uint64_t* p_full = s_full + 2;                  // [2]  softmax -> MMA: P_t(j) is in smem, S_t and O_t are released
  uint64_t* o_done = p_full + 2;                  // [2]  MMA -> softmax: the last PV_t has completed
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(o_done + 2);

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
    }
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
    if (lane == 0) {
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
    if (lane == 0) {
      // ------------------------------- MMA issuer ---------------------------------
      constexpr uint32_t IDESC_S = umma_idesc_bf16(128, 128, 0, 0);
      constexpr uint32_t IDESC_O = umma_idesc_bf16(128, 64, 0, 1);  // B = V is MN-major
      // descriptors: only the 14-bit start-address field changes; built once, advanced by adds (single issuing lane)
      const uint64_t dK = umma_desc_sw128(0, 16, 1024);    // K-major operands (Q, K_j, P)
      const uint64_t dV = umma_desc_sw128(0, 1024, 1024);  // MN-major operand (V_j)
      const uint32_t q_lo = smem_u32(sQ) >> 4, k_lo = smem_u32(sK) >> 4, v_lo = smem_u32(sV) >> 4;
      auto issue_s = [&](int t, int st) {
        const uint64_t qd = dK | (uint64_t)(q_lo + t * (AT_TILE_BYTES >> 4));
        const uint64_t kd = dK | (uint64_t)(k_lo + st * (AT_TILE_BYTES >> 4));
        umma_ss(tmem_base + t * 128, qd, kd, IDESC_S, 0u);
#pragma unroll
        for (int k = 1; k < AT_D / 16; ++k) umma_ss(tmem_base + t * 128, qd + 2 * k, kd + 2 * k, IDESC_S, 1u);
        umma_commit(&s_full[t]);
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
        for (int t = 0; t < 2; ++t) {
          AT_STAMP(2, j, t * 2);
          mbar_wait(&p_full[t], j & 1);
          tc_fence_after();
          AT_STAMP(2, j, t * 2 + 1);
          const uint64_t vd = dV | (uint64_t)(v_lo + st * (AT_TILE_BYTES >> 4));
#pragma unroll
          for (int k = 0; k < AT_BKV / 16; ++k)  // A = P_t in TMEM over S_t: 8 columns (16 keys) per k-step; V: 16 rows
            umma_ts(tmem_base + 256 + t * 64, tmem_base + t * 128 + 8 * k, vd + 128 * k, IDESC_O, (j | k) != 0 ? 1u : 0u);
          if (j + 1 < nkv) issue_s(t, stn);  // S_t(j+1): its commit also covers PV_t(j)
          else umma_commit(&o_done[t]);
        }
        umma_commit(&kv_empty[st]);  // K_j / V_j have been read by every MMA issued so far
      }
    }
  } else {
    // ------------------------------- softmax / output ---------------------------
    const int t = warp >> 2;                 // query tile of this warp group
    const int row = (warp & 3) * 32 + lane;  // row inside the tile = TMEM lane
    const uint32_t lane_off = static_cast<uint32_t>((warp & 3) * 32) << 16;
    const uint32_t tS = tmem_base + t * 128 + lane_off;
    const uint32_t tO = tmem_base + 256 + t * 64 + lane_off;
    float m_ref = -INFINITY, l = 0.0f;
    // Started together the two warp groups run in lock step: both exponentiate at once (contending for the MUFU) and
    // both idle while the MMAs run.  Tile 1 therefore begins about half a block period late; the offset sustains
    // itself because each group's S / P*V round trip then falls into the other group's exponentiation phase.
    if (t == 1 && nkv > 1) __nanosleep(AT_STAGGER_NS);

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
      if (j == 0) {
        m_ref = bm;
      } else {
        const bool need = (bm - m_ref) * scale_log2 > AT_LAZY;
        if (__any_sync(0xffffffffu, need)) {
          // raise the reference maximum of the rows that need it and rescale their O accumulators in TMEM
          // (PV_t(j-1) is complete: the commit that raised s_full covers it)
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
      AT_STAMP(t, j, 2);
      const float moff = m_ref * scale_log2;
      float ps4[4] = {0.0f, 0.0f, 0.0f, 0.0f};
#pragma unroll
      for (int i = 0; i < AT_BKV; i += 4) {
#pragma unroll
        for (int q = 0; q < 4; ++q) {
          s[i + q] = fast_exp2(fmaf(s[i + q], scale_log2, -moff));  // exp2(-inf) = 0 for padded keys
          ps4[q] += s[i + q];
        }
      }
      l += (ps4[0] + ps4[1]) + (ps4[2] + ps4[3]);
      // P_t(j) as bf16 pairs into the first 64 columns of the S_t region (S is in registers by now): the P*V MMA reads
      // its A operand straight from tensor memory -- no shared-memory round trip and no operand bandwidth for P.  The
      // next S_t overwrites the region only after P*V (MMAs of one thread execute in issue order).
      {
        uint32_t* pk = reinterpret_cast<uint32_t*>(s);
#pragma unroll
        for (int i = 0; i < AT_BKV / 2; ++i) pk[i] = pack_bf16x2(s[2 * i], s[2 * i + 1]);
        tmem_st_32x32(tS, s);
        tmem_st_32x32(tS + 32, s + 32);
        tmem_st_wait();
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&p_full[t]);
      AT_STAMP(t, j, 3);
    }

    if (nkv > 0) mbar_wait(&o_done[t], 0);  // (no keys at all: O is never produced, the rows below store 0 * inf = NaN
    tc_fence_after();                       //  like softmax over an empty set in the reference)
    const int r = q0 + t * AT_BQ + row;
    const float inv = 1.0f / l;
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

void tc_attention(cudaStream_t st, const bf16* qkv, bf16* out, int batches, int rows, int heads, const int* kv_len,
                  float scale, int out_f16) {
  const int inner = heads * AT_D;
  CUtensorMap tm;
  uint64_t dims[3] = {(uint64_t)3 * inner, (uint64_t)rows, (uint64_t)batches};
  uint64_t strides[2] = {(uint64_t)3 * inner * 2, (uint64_t)3 * inner * 2 * (uint64_t)rows};
  uint32_t box[3] = {AT_D, 128, 1};
  make_tmap_bf16(&tm, qkv, 3, dims, strides, box);
  dim3 grid((rows + 2 * AT_BQ - 1) / (2 * AT_BQ), heads, batches);
  launch_pdl_id(2, tc_attn_kernel, grid, dim3(320), AT_SMEM, st, tm, out, rows, heads, kv_len,
             scale * 1.4426950408889634f, out_f16);
  F5B_POST_LAUNCH();
}

void attn_trace_enable(int on) { F5B_CHECK_CUDA(cudaMemcpyToSymbol(g_attn_trace_on, &on, sizeof(int))); }
void attn_trace_read(long long* host) {
  F5B_CHECK_CUDA(cudaMemcpyFromSymbol(host, g_attn_trace, sizeof(long long) * 3 * 64 * 4));
}

void tc_init_kernels() {
  F5B_CHECK_CUDA(cudaFuncSetAttribute(tc_attn_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, AT_SMEM));
}

}  // namespace f5b
