// Dense contraction on the 5th-gen tensor cores: C[M,N] = A[M,K] * W[N,K]^T (+ fused epilogue).
//
// Replaces every nn.Linear on the reference hot path (reference src/f5_tts/model/modules.py:312,333,355-356,
// 393-395,418,781; backbones/dit.py:124,205,208; backbones/unett.py:92,171,186) and, in "conv" mode, the grouped
// k=31 Conv1d of ConvPositionEmbedding (modules.py:174-179) as an implicit GEMM whose A tiles are TMA loads at
// shifted row coordinates (zero fill outside the sequence comes from the TMA out-of-bounds rule).
//
// Structure (one CTA per SM, persistent over a static tile schedule):
//   warp 0 lane 0 : TMA producer   -- cp.async.bulk.tensor into a STAGES-deep 128B-swizzled smem ring
//   warp 1 lane 0 : MMA issuer     -- tcgen05.mma.cta_group::1.kind::f16, 128 x BN x 16 per instruction,
//                                     fp32 accumulators in TMEM, double buffered (2 x BN columns)
//   warps 2..9    : epilogue       -- two groups of four warps (a warp reads the TMEM lane quarter warp%4); group g owns
//                                     the 64-column units g, g+2, ... of the tile: tcgen05.ld 32x32b, fused bias / RoPE /
//                                     activation / gate, swizzled smem staging, TMA store or L2 reduce-add; overlaps the
//                                     next tile's MMAs.  Two warps per scheduler: the in-situ timeline showed the
//                                     4-warp epilogue exposed for 5.5 us on the 1-tile-per-CTA shapes
#include "common.cuh"
#include "sm100_ptx.cuh"
#include "tc_gemm.h"

#include <cudaTypedefs.h>
#include <cstdlib>
#include <mutex>

namespace f5b {

using namespace ptx;

static constexpr int BM = 128;
static constexpr int BK = 64;  // 64 bf16 = 128 B = one swizzle row
static constexpr int A_STAGE_BYTES = BM * BK * 2;

// ---------------------------------------------------------------------------------------------
// epilogue: TMEM -> registers (bias / RoPE / activation / gate) -> 128B-swizzled smem staging -> TMA store, or TMA
// reduce-add (fp32, performed in L2) for the gate * y + residual update.  No per-thread global stores: every global
// write is a full-line bulk transfer, and the residual is never read back by the SM.
// ---------------------------------------------------------------------------------------------
// Epilogue variants.  The generic epilogue (any activation / output type / RoPE / gate, chosen at run time) is ~4600
// SASS instructions per chunk iteration of which ~500 execute: ncu's source view showed the epilogue warps mostly
// stalled on instruction fetch (no_inst) while hopping over the unused cases.  The three shapes of the per-step chain get
// their own instantiation whose chunk loop holds only the code it runs; everything else uses the generic one.
enum { EV_GENERIC = 0, EV_STORE16 = 1, EV_RESID = 2, EV_GELU16 = 3 };
//   EV_STORE16: no activation, 16-bit store, optional RoPE, no gate          (qkv, plain projections)
//   EV_RESID  : no activation, fp32 L2 reduce-add, optional per-column gate  (attention out-projection, ff2)
//   EV_GELU16 : tanh-GELU, 16-bit store, no RoPE, no gate                    (ff1)

static constexpr int EPI_GROUPS = 2;  // epilogue warp groups
static constexpr int EPI_NBUF = 2;    // staging buffers per group
static constexpr int EPI_BUF_BYTES = BM * 128;  // 128 rows x 128 B (32 fp32 or 64 16-bit columns)

// Activations of the tensor-core epilogue: single-MUFU forms (tanh.approx, ex2.approx, rcp.approx).  Their error
// (~2^-11 relative) is below the rounding of the 16-bit value that is stored; the fp32 mode keeps libm (common.cuh).
__device__ __forceinline__ float fast_tanh(float x) {
  float y;
  asm("tanh.approx.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}
template <int ACT>
__device__ __forceinline__ float act_fast(float x) {
  if (ACT == ACT_GELU_TANH) {
    const float u = 0.7978845608028654f * (x + 0.044715f * x * x * x);
    return 0.5f * x * (1.0f + fast_tanh(u));
  } else if (ACT == ACT_MISH) {
    // x * tanh(softplus(x)) = x * (n^2 - 1) / (n^2 + 1), n = 1 + e^x
    const float e = __expf(fminf(x, 20.0f));
    const float n2 = (1.0f + e) * (1.0f + e);
    return x * __fdividef(n2 - 1.0f, n2 + 1.0f);
  } else if (ACT == ACT_SILU) {
    return __fdividef(x, 1.0f + __expf(-x));
  } else if (ACT == ACT_GELU_ERF) {
    return act_apply(ACT_GELU_ERF, x);  // exact erf: ConvNeXt / Vocos blocks, off the per-step path
  }
  return x;
}

// bias_s / gate_s: this tile's bias and gate columns staged in shared memory (loaded while the MMAs run);
// v[0..32) are accumulator columns [n, n+32) of one row, cs = offset of the chunk inside the tile.
template <int HALF>
__device__ __forceinline__ void rope_rotate(float* v, const float4 (&t)[16]) {
#pragma unroll
  for (int j = 0; j < 32; j += 4) {
    const float4 tt = t[HALF * 8 + (j >> 2)];  // (cos, sin) of two consecutive interleaved pairs
    float a = v[j], c = v[j + 1], e = v[j + 2], f = v[j + 3];
    v[j] = a * tt.x - c * tt.y;
    v[j + 1] = c * tt.x + a * tt.y;
    v[j + 2] = e * tt.z - f * tt.w;
    v[j + 3] = f * tt.z + e * tt.w;
  }
}

// ROPE: the variant can carry rotary columns; GATE: 0 = never, 1 = staged per-column gate only, 2 = staged or per-row.
template <int ACT, bool ROPE, int GATE>
__device__ __forceinline__ void epi_transform(const EpiParams& p, float* v, int n, int cs, int b, int pos, bool dead,
                                              const float* bias_s, const float* gate_s, const float4 (&rope)[16],
                                              bool has_rope) {
#pragma unroll
  for (int j = 0; j < 32; j += 4) {
    const float4 bv = *reinterpret_cast<const float4*>(bias_s + cs + j);
    v[j] += bv.x; v[j + 1] += bv.y; v[j + 2] += bv.z; v[j + 3] += bv.w;
  }
  if (ROPE && has_rope && n < p.rope_cols) {
    const int within = n % p.rope_period;
    if ((within >> 6) < p.rope_heads) {
      // rope[0..16): this row's 32 (cos, sin) pairs, loaded once per tile; a 32-column chunk is half a head
      if (within & 32) rope_rotate<1>(v, rope);
      else rope_rotate<0>(v, rope);
    }
  }
  if (ACT != ACT_NONE) {
#pragma unroll
    for (int j = 0; j < 32; ++j) v[j] = act_fast<ACT>(v[j]);
  }
  if (GATE != 0 && p.gate != nullptr) {
    if (GATE == 1 || p.gate_bmod <= 1) {
#pragma unroll
      for (int j = 0; j < 32; j += 4) {
        const float4 gv = *reinterpret_cast<const float4*>(gate_s + cs + j);
        v[j] *= gv.x; v[j + 1] *= gv.y; v[j + 2] *= gv.z; v[j + 3] *= gv.w;
      }
    } else {  // per-batch-element gates (f5b_forward with different times per utterance): row dependent, read directly
      const float* g = p.gate + (long long)(b % p.gate_bmod) * p.gate_bstride + n;
#pragma unroll
      for (int j = 0; j < 32; ++j)
        if (n + j < p.n_valid) v[j] *= __ldg(g + j);
    }
  }
  if (dead) {
#pragma unroll
    for (int j = 0; j < 32; ++j) v[j] = 0.0f;
  }
}

template <int BN, int STAGES, int VAR, int PAIR>
__global__ void __launch_bounds__(64 + 128 * EPI_GROUPS, 1)
tc_gemm_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
               const __grid_constant__ CUtensorMap tmO, const TcGemmShape s, const EpiParams ep) {
  // PAIR: two CTAs (one TPC) share a 256-row tile -- each keeps its own 128 A rows and half of the B tile, so a CTA
  // takes in A + B/2 per k-block instead of A + B: the mainloop is bound by the ~64 B/clk an SM can pull from L2
  // (0.43 us per 48 KB k-block at BN = 256), not by the tensor pipe.
  constexpr int BNL = PAIR ? BN / 2 : BN;  // B rows in this CTA's shared memory
  constexpr int B_STAGE_BYTES = BNL * BK * 2;
  constexpr uint32_t STAGE_TX = (A_STAGE_BYTES + B_STAGE_BYTES) * (PAIR ? 2 : 1);  // counted on the leader's barrier
  constexpr uint32_t ACC_COLS = 2 * BN <= 256 ? 2 * BN : 512;  // power of two; the two accumulators sit at 0 and BN

  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
  uint8_t* sA = smem;
  uint8_t* sB = smem + STAGES * A_STAGE_BYTES;
  uint8_t* sO = smem + STAGES * (A_STAGE_BYTES + B_STAGE_BYTES);
  float* sBias = reinterpret_cast<float*>(sO + EPI_GROUPS * EPI_NBUF * EPI_BUF_BYTES);  // [BN] bias, [BN] gate of the current tile
  float* sGate = sBias + BN;
  uint64_t* full = reinterpret_cast<uint64_t*>(sGate + BN);
  uint64_t* empty = full + STAGES;
  uint64_t* tfull = empty + STAGES;
  uint64_t* tempty = tfull + 2;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tempty + 2);

  KT_DECL
  KT_BEGIN();
  pdl_trigger();
  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const uint32_t rank = PAIR ? cluster_ctarank() : 0u;  // 0 = leader: issues the MMAs, owns full[] and tempty[]
  const int mrows = PAIR ? (s.tiles_m_per_batch + 1) / 2 : s.tiles_m_per_batch;  // tile rows (of BM * (1 + PAIR)) per batch
  const int total_tiles = mrows * s.batches * s.tiles_n * s.splits;  // work units: (tile, k-split)
  const int unit0 = PAIR ? (int)(blockIdx.x >> 1) : (int)blockIdx.x;
  const int unit_step = PAIR ? (int)(gridDim.x >> 1) : (int)gridDim.x;

  if (threadIdx.x == 0) {
    prefetch_tmap(&tmA);
    prefetch_tmap(&tmB);
    prefetch_tmap(&tmO);
    for (int i = 0; i < STAGES; ++i) {
      mbar_init(&full[i], 1);
      mbar_init(&empty[i], 1);
    }
    for (int i = 0; i < 2; ++i) {
      mbar_init(&tfull[i], 1);
      mbar_init(&tempty[i], 4 * EPI_GROUPS * (PAIR ? 2 : 1));
    }
    fence_barrier_init();
  }
  if (warp == 1) {
    if (PAIR) tmem_alloc_pair(tmem_slot, ACC_COLS);
    else tmem_alloc(tmem_slot, ACC_COLS);
  }
  tc_fence_before();
  if (PAIR) cluster_sync_all();  // the peer's barriers are initialised before anything is signalled on them
  else __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  // Single-lane roles are entered through elect.sync, not a lane test: the compiler then knows exactly one lane is active
  // and feeds tcgen05.mma / cp.async.bulk.tensor straight from uniform registers.  Behind `lane == 0` it wraps EVERY such
  // instruction in a loop over the possibly-active lanes (ELECT / R2UR / BRA.U.ANY): 257 of those loops in this file
  // before the change, ~70 cycles of issue per MMA against a 64-cycle tensor floor at BN = 128.
  const bool lead_warp0 = warp == 0, lead_warp1 = warp == 1 && rank == 0;
  if (lead_warp0 && elect_one()) {
    // ------------------------------- TMA producer -------------------------------
    // The weights do not depend on the previous kernel: the W tiles of the first ring fill are requested BEFORE the
    // programmatic-dependency wait, so their HBM latency overlaps the predecessor's tail; the A tiles follow after it.
    uint32_t it = 0;
    int pre = 0;
    const uint32_t full0 = PAIR ? cluster_map(&full[0], 0) : 0u;  // the leader's full[] barriers count both CTAs' bytes
    if (unit0 < total_tiles) {
      const int ks = unit0 % s.splits, tile = unit0 / s.splits;
      const int tn = tile % s.tiles_n;
      const int kb0 = ks * s.num_kb / s.splits, kb1 = (ks + 1) * s.num_kb / s.splits;
      pre = kb1 - kb0 < STAGES ? kb1 - kb0 : STAGES;
      for (int i = 0; i < pre; ++i) {
        if (rank == 0) mbar_expect_tx(&full[i], STAGE_TX);
        if (PAIR) tma_load_2d_pair(sB + i * B_STAGE_BYTES, &tmB, full0 + 8 * i, (kb0 + i) * BK, tn * BN + (int)rank * BNL);
        else tma_load_2d(sB + i * B_STAGE_BYTES, &tmB, &full[i], (kb0 + i) * BK, tn * BN);
      }
    }
    pdl_wait();
    KT_WAITED();
    for (int unit = unit0; unit < total_tiles; unit += unit_step) {
      const int ks = unit % s.splits, tile = unit / s.splits;
      const int tm = tile / s.tiles_n, tn = tile - tm * s.tiles_n;
      const int bz = tm / mrows;
      const int r0 = ((tm - bz * mrows) * (PAIR ? 2 : 1) + (int)rank) * BM;  // beyond the batch for an odd tail: zero fill
      const int n0 = tn * BN;
      const int kb0 = ks * s.num_kb / s.splits, kb1 = (ks + 1) * s.num_kb / s.splits;
      for (int kb = kb0; kb < kb1; ++kb, ++it) {
        const uint32_t st = it % STAGES, ph = (it / STAGES) & 1u;
        const bool w_requested = (int)it < pre;  // first ring fill: W is already in flight, the slot is fresh
        if (!w_requested) {
          mbar_wait(&empty[st], ph ^ 1u);
          if (rank == 0) mbar_expect_tx(&full[st], STAGE_TX);
        }
        int a_c0 = kb * BK, a_c1 = r0;
        if (s.conv) {
          const int tap = kb / s.conv_kb_per_tap, sub = kb - tap * s.conv_kb_per_tap;
          a_c0 = (n0 / s.conv_cg) * s.conv_cg + sub * BK;
          a_c1 = r0 + tap - s.conv_pad;
        }
        if (PAIR) {
          tma_load_3d_pair(sA + st * A_STAGE_BYTES, &tmA, full0 + 8 * st, a_c0, a_c1, bz);
          if (!w_requested) tma_load_2d_pair(sB + st * B_STAGE_BYTES, &tmB, full0 + 8 * st, kb * BK, n0 + (int)rank * BNL);
        } else {
          tma_load_3d(sA + st * A_STAGE_BYTES, &tmA, &full[st], a_c0, a_c1, bz);
          if (!w_requested) tma_load_2d(sB + st * B_STAGE_BYTES, &tmB, &full[st], kb * BK, n0);
        }
      }
    }
  } else if (lead_warp1 && elect_one()) {
    // ------------------------------- MMA issuer ---------------------------------
    // Descriptors are built once: only the 14-bit start-address field (bytes >> 4) changes between MMAs, and it cannot
    // carry out of its field (smem addresses < 256 KB), so a stage / k-step advance is a plain add.  The issuing thread
    // is a single lane: every instruction it spends per MMA is serial latency on the tensor pipe's critical path.
    const uint64_t desc0 = umma_desc_sw128(0, 16, 1024);
    const uint32_t a_lo = (smem_u32(sA) & 0x3FFFFu) >> 4, b_lo = (smem_u32(sB) & 0x3FFFFu) >> 4;
    const uint32_t idesc = s.idesc;
    uint32_t it = 0, tcount = 0;
    for (int unit = unit0; unit < total_tiles; unit += unit_step, ++tcount) {
      const int ks = unit % s.splits;
      const int kb0 = ks * s.num_kb / s.splits, kb1 = (ks + 1) * s.num_kb / s.splits;
      const uint32_t acc = tcount & 1u, aph = (tcount >> 1) & 1u;
      mbar_wait(&tempty[acc], aph ^ 1u);
      tc_fence_after();
      const uint32_t d_tmem = tmem_base + acc * BN;
      uint32_t st = it % STAGES, ph = (it / STAGES) & 1u;
      for (int kb = kb0; kb < kb1; ++kb, ++it) {
        mbar_wait(&full[st], ph);
        if (it == 0) KT_MARK(2);  // first operands have landed
        tc_fence_after();
        const uint64_t ad = desc0 | (uint64_t)(a_lo + st * (A_STAGE_BYTES >> 4));
        const uint64_t bd = desc0 | (uint64_t)(b_lo + st * (B_STAGE_BYTES >> 4));
        if (PAIR) {
          umma_ss_pair(d_tmem, ad, bd, idesc, kb != kb0 ? 1u : 0u);
#pragma unroll
          for (int k = 1; k < BK / 16; ++k) umma_ss_pair(d_tmem, ad + 2 * k, bd + 2 * k, idesc, 1u);
          umma_commit_pair(&empty[st]);  // frees the slot in both CTAs
        } else {
          umma_ss(d_tmem, ad, bd, idesc, kb != kb0 ? 1u : 0u);
#pragma unroll
          for (int k = 1; k < BK / 16; ++k) umma_ss(d_tmem, ad + 2 * k, bd + 2 * k, idesc, 1u);
          umma_commit(&empty[st]);  // frees the smem slot once these MMAs have read it
        }
        if (++st == STAGES) {
          st = 0;
          ph ^= 1u;
        }
      }
      if (PAIR) umma_commit_pair(&tfull[acc]);  // accumulator complete (each CTA's epilogue waits on its own barrier)
      else umma_commit(&tfull[acc]);
      KT_MARK(3);  // (last tile wins) all MMAs issued
    }
  } else if (warp >= 2) {
    // ------------------------------- epilogue -----------------------------------
    const int quarter = warp & 3;  // TMEM lanes [32*quarter, 32*quarter+32) are this warp's
    const int rl = quarter * 32 + lane;  // row inside the tile
    const int grp = (warp - 2) >> 2;  // units of 64 columns alternate between the two groups
    const bool issuer_warp = ((warp - 2) & 3) == 0;  // its elected lane (deterministic for a full mask) issues the bulk stores
    const bool issuer = issuer_warp && lane == 0;      // trace stamps only
    uint8_t* sOg = sO + grp * EPI_NBUF * EPI_BUF_BYTES;
    constexpr bool GEN = VAR == EV_GENERIC;
    constexpr bool MAY_ROPE = GEN || VAR == EV_STORE16;
    const bool out32 = GEN ? (ep.kind == EPI_RESID || ep.out_type == OUT_F32) : (VAR == EV_RESID);
    const bool resid = GEN ? ep.kind == EPI_RESID : (VAR == EV_RESID);
    uint32_t tcount = 0, nstore = 0;
    // Dry pass: while the first tile's MMAs run, the epilogue warps walk one 64-column unit of the epilogue on whatever
    // the accumulator holds -- no barrier wait, no bulk store -- so the code is in the instruction cache when the tile
    // completes (the first chunk of a launch used to cost 2-3x a warm one, exposed on the one-tile-per-CTA shapes).
    bool dry = true;
    const uint32_t tempty0 = PAIR ? cluster_map(&tempty[0], 0) : 0u;
    for (int unit = unit0; unit < total_tiles;) {
      const int ks = unit % s.splits, tile = unit / s.splits;
      const int tm = tile / s.tiles_n, tn = tile - tm * s.tiles_n;
      const int bz = tm / mrows;
      const int r0 = ((tm - bz * mrows) * (PAIR ? 2 : 1) + (int)rank) * BM;
      const int n0 = tn * BN;
      const uint32_t acc = tcount & 1u, aph = (tcount >> 1) & 1u;
      const bool with_bias = ks == 0;  // the bias joins exactly one of the partial sums
      const int r = r0 + rl;
      const bool row_ok = r < s.rows_per_batch;
      const long long m = (long long)bz * s.rows_per_batch + r;
      int b = 0, pos = 0;
      bool dead = !row_ok;
      if (row_ok) {
        b = (int)(m / ep.rows_per_batch);
        pos = (int)(m - (long long)b * ep.rows_per_batch);
        if (ep.row_len != nullptr && pos >= ep.row_len[b]) dead = true;
      }
      // stage this tile's bias / gate columns while its MMAs are still running
      for (int i = rl; i < BN; i += 128) {
        if (((i >> 6) & 1) != grp) continue;  // each group stages (and later reads) only its own columns
        const int n = n0 + i;
        const bool in = n < ep.n_valid;
        sBias[i] = (ep.bias != nullptr && with_bias && in) ? __ldg(ep.bias + n) : 0.0f;
        if (GEN || VAR == EV_RESID) sGate[i] = (ep.gate != nullptr && ep.gate_bmod <= 1 && in) ? __ldg(ep.gate + n) : 1.0f;
      }
      named_bar_sync(1 + grp, 128);
      float4 rope_reg[16];
      const bool has_rope = MAY_ROPE && ep.rope_cs != nullptr && n0 < ep.rope_cols;  // uniform per tile
      if (!MAY_ROPE) {
      } else if (has_rope) {
        const float4* src = reinterpret_cast<const float4*>(ep.rope_cs + (long long)pos * 64);
#pragma unroll
        for (int i = 0; i < 16; ++i) rope_reg[i] = __ldg(src + i);
      } else {
#pragma unroll
        for (int i = 0; i < 16; ++i) rope_reg[i] = make_float4(1.0f, 0.0f, 1.0f, 0.0f);
      }
      if (!dry) mbar_wait(&tfull[acc], aph);
      if (issuer && grp == 0) KT_MARK(4);  // (last tile wins) accumulator complete, epilogue starts
      tc_fence_after();
      const uint32_t t_row = tmem_base + acc * BN + (static_cast<uint32_t>(quarter * 32) << 16);
#pragma unroll 1
      for (int c = grp * 64; c < BN; c += (c & 32) ? 96 : 32) {
        const int n = n0 + c;
        if (n >= s.N) break;  // uniform: the rest of the tile lies beyond the last column
        float v[32];
        tmem_ld_32x32(t_row + c, v);
        tmem_ld_wait();
        if (GEN) {
          switch (ep.act) {
            case ACT_GELU_TANH: epi_transform<ACT_GELU_TANH, true, 2>(ep, v, n, c, b, pos, dead, sBias, sGate, rope_reg, has_rope); break;
            case ACT_GELU_ERF: epi_transform<ACT_GELU_ERF, true, 2>(ep, v, n, c, b, pos, dead, sBias, sGate, rope_reg, has_rope); break;
            case ACT_MISH: epi_transform<ACT_MISH, true, 2>(ep, v, n, c, b, pos, dead, sBias, sGate, rope_reg, has_rope); break;
            case ACT_SILU: epi_transform<ACT_SILU, true, 2>(ep, v, n, c, b, pos, dead, sBias, sGate, rope_reg, has_rope); break;
            default: epi_transform<ACT_NONE, true, 2>(ep, v, n, c, b, pos, dead, sBias, sGate, rope_reg, has_rope); break;
          }
        } else if (VAR == EV_STORE16) {
          epi_transform<ACT_NONE, true, 0>(ep, v, n, c, b, pos, dead, sBias, sGate, rope_reg, has_rope);
        } else if (VAR == EV_RESID) {
          epi_transform<ACT_NONE, false, 1>(ep, v, n, c, b, pos, dead, sBias, sGate, rope_reg, false);
        } else {
          epi_transform<ACT_GELU_TANH, false, 0>(ep, v, n, c, b, pos, dead, sBias, sGate, rope_reg, false);
        }
        uint8_t* buf = sOg + (nstore % EPI_NBUF) * EPI_BUF_BYTES;
        uint8_t* row = buf + rl * 128;
        const int sw = rl & 7;
        bool flush;
        int c_first;
        if (out32) {
#pragma unroll
          for (int j = 0; j < 8; ++j)
            *reinterpret_cast<float4*>(row + ((j ^ sw) << 4)) =
                make_float4(v[4 * j], v[4 * j + 1], v[4 * j + 2], v[4 * j + 3]);
          flush = true;
          c_first = n;
        } else {
          const int half = (c >> 5) & 1;
          const bool f16 = ep.out_type == OUT_F16;
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            uint4 pk;
            if (f16) {
              pk.x = pack_f16x2(v[8 * j], v[8 * j + 1]);
              pk.y = pack_f16x2(v[8 * j + 2], v[8 * j + 3]);
              pk.z = pack_f16x2(v[8 * j + 4], v[8 * j + 5]);
              pk.w = pack_f16x2(v[8 * j + 6], v[8 * j + 7]);
            } else {
              pk.x = pack_bf16x2(v[8 * j], v[8 * j + 1]);
              pk.y = pack_bf16x2(v[8 * j + 2], v[8 * j + 3]);
              pk.z = pack_bf16x2(v[8 * j + 4], v[8 * j + 5]);
              pk.w = pack_bf16x2(v[8 * j + 6], v[8 * j + 7]);
            }
            *reinterpret_cast<uint4*>(row + (((half * 4 + j) ^ sw) << 4)) = pk;
          }
          flush = half == 1 || c + 32 >= BN || n + 32 >= s.N;
          c_first = n - half * 32;
        }
        if (flush) {
          fence_proxy_async_smem();  // generic-proxy smem writes -> visible to the TMA engine
          // two staging buffers per group: the store issued at the previous flush has to be done reading its buffer
          // before anyone passes this barrier and starts to fill it again (it had a whole chunk of time to do so)
          if (issuer_warp && !dry && elect_one()) bulk_wait_read<0>();
          named_bar_sync(1 + grp, 128);
          if (dry) break;
          if (issuer_warp && r0 < s.rows_per_batch && elect_one()) {
            if (resid) tma_reduce_add_3d(&tmO, buf, c_first, r0 - ep.row_skip, bz);
            else tma_store_3d(&tmO, buf, c_first, r0 - ep.row_skip, bz);
            bulk_commit();
          }
          ++nstore;
        }
      }
      if (dry) {
        dry = false;
        continue;  // now the same unit for real
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) {
        if (PAIR) mbar_arrive_cluster(tempty0 + 8 * acc);
        else mbar_arrive(&tempty[acc]);
      }
      unit += unit_step;
      ++tcount;
    }
    if (issuer && grp == 0) KT_MARK(5);  // last store issued
    // the staging buffers must outlive the reads of the bulk stores; the writes themselves complete with the grid
    if (issuer_warp && elect_one()) bulk_wait_read<0>();
  }

  tc_fence_before();
  if (PAIR) cluster_sync_all();  // neither CTA may leave while its peer's MMAs / barrier arrivals still touch it
  else __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    if (PAIR) tmem_dealloc_pair(tmem_base, ACC_COLS);
    else tmem_dealloc(tmem_base, ACC_COLS);
  }
  KT_END(1);
}

KT_DEFINE_SETTER(ktrace_set_tc_gemm)

// ---------------------------------------------------------------------------------------------
// host side
// ---------------------------------------------------------------------------------------------
static PFN_cuTensorMapEncodeTiled_v12000 g_encode = nullptr;
static std::once_flag g_encode_once;

static void load_encode() {
  std::call_once(g_encode_once, [] {
    void* fn = nullptr;
    cudaDriverEntryPointQueryResult q;
    cudaError_t e = cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &q);
    if (e == cudaSuccess && q == cudaDriverEntryPointSuccess) g_encode = reinterpret_cast<PFN_cuTensorMapEncodeTiled_v12000>(fn);
  });
}

void make_tmap_bf16(CUtensorMap* out, const void* base, int rank, const uint64_t* dims, const uint64_t* strides_bytes,
                    const uint32_t* box, int fp16) {
  make_tmap(out, base, rank, dims, strides_bytes, box, fp16 ? OUT_F16 : OUT_BF16);
}

void make_tmap(CUtensorMap* out, const void* base, int rank, const uint64_t* dims, const uint64_t* strides_bytes,
               const uint32_t* box, int elem_type) {
  load_encode();
  F5B_REQUIRE(g_encode != nullptr, "cuTensorMapEncodeTiled unavailable (no CUDA driver)");
  F5B_REQUIRE((reinterpret_cast<uintptr_t>(base) & 15) == 0, "TMA base must be 16-byte aligned");
  cuuint64_t gd[5];
  cuuint64_t gs[5];
  cuuint32_t bx[5], es[5];
  for (int i = 0; i < rank; ++i) {
    gd[i] = dims[i];
    bx[i] = box[i];
    es[i] = 1;
  }
  for (int i = 0; i + 1 < rank; ++i) {
    gs[i] = strides_bytes[i];
    F5B_REQUIRE(gs[i] % 16 == 0, "TMA strides must be multiples of 16 bytes");
  }
  const CUtensorMapDataType dt = elem_type == OUT_F32 ? CU_TENSOR_MAP_DATA_TYPE_FLOAT32
                                 : elem_type == OUT_F16 ? CU_TENSOR_MAP_DATA_TYPE_FLOAT16
                                                        : CU_TENSOR_MAP_DATA_TYPE_BFLOAT16;
  CUresult r = g_encode(out, dt, (cuuint32_t)rank, const_cast<void*>(base), gd, gs, bx,
                        es, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                        CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    set_error("cuTensorMapEncodeTiled failed with CUresult " + std::to_string((int)r));
    throw Err{F5B_ERR_CUDA};
  }
}

template <int BN, int STAGES, int PAIR>
static constexpr int smem_bytes() {
  return STAGES * (A_STAGE_BYTES + (PAIR ? BN / 2 : BN) * BK * 2) + EPI_GROUPS * EPI_NBUF * EPI_BUF_BYTES + 2 * BN * 4 +
         1024 + 256;
}
static int epilogue_variant(const EpiParams& ep) {
  if (ep.kind == EPI_STORE && ep.out_type != OUT_F32 && ep.gate == nullptr) {
    if (ep.act == ACT_NONE) return EV_STORE16;
    if (ep.act == ACT_GELU_TANH && ep.rope_cs == nullptr) return EV_GELU16;
  }
  if (ep.kind == EPI_RESID && ep.act == ACT_NONE && ep.rope_cs == nullptr && (ep.gate == nullptr || ep.gate_bmod <= 1))
    return EV_RESID;
  return EV_GENERIC;
}

template <int BN, int STAGES, int VAR, int PAIR>
static void launch_var(cudaStream_t st, int grid, const CUtensorMap& ta, const CUtensorMap& tb, const CUtensorMap& to,
                       const TcGemmShape& s, const EpiParams& ep) {
  launch_pdl_cluster(1, PAIR ? 2 : 1, tc_gemm_kernel<BN, STAGES, VAR, PAIR>, dim3(grid), dim3(64 + 128 * EPI_GROUPS),
                     smem_bytes<BN, STAGES, PAIR>(), st, ta, tb, to, s, ep);
  F5B_POST_LAUNCH();
}
template <int BN, int STAGES, int PAIR>
static void launch_cfg(cudaStream_t st, int grid, const CUtensorMap& ta, const CUtensorMap& tb, const CUtensorMap& to,
                       const TcGemmShape& s, const EpiParams& ep) {
  constexpr bool SPECIAL = BN >= 128;  // the narrow tile only serves the conv / proj_out launches: generic epilogue
  switch (epilogue_variant(ep)) {
    case EV_STORE16: return launch_var<BN, STAGES, SPECIAL ? EV_STORE16 : EV_GENERIC, PAIR>(st, grid, ta, tb, to, s, ep);
    case EV_RESID: return launch_var<BN, STAGES, SPECIAL ? EV_RESID : EV_GENERIC, PAIR>(st, grid, ta, tb, to, s, ep);
    case EV_GELU16: return launch_var<BN, STAGES, SPECIAL ? EV_GELU16 : EV_GENERIC, PAIR>(st, grid, ta, tb, to, s, ep);
    default: return launch_var<BN, STAGES, EV_GENERIC, PAIR>(st, grid, ta, tb, to, s, ep);
  }
}

template <int BN, int STAGES, int PAIR>
static void config_one() {
  constexpr bool SPECIAL = BN >= 128;
  constexpr int bytes = smem_bytes<BN, STAGES, PAIR>();
  static_assert(bytes <= 232448, "shared memory budget");
  F5B_CHECK_CUDA(cudaFuncSetAttribute(tc_gemm_kernel<BN, STAGES, EV_GENERIC, PAIR>,
                                      cudaFuncAttributeMaxDynamicSharedMemorySize, bytes));
  F5B_CHECK_CUDA(cudaFuncSetAttribute(tc_gemm_kernel<BN, STAGES, SPECIAL ? EV_STORE16 : EV_GENERIC, PAIR>,
                                      cudaFuncAttributeMaxDynamicSharedMemorySize, bytes));
  F5B_CHECK_CUDA(cudaFuncSetAttribute(tc_gemm_kernel<BN, STAGES, SPECIAL ? EV_RESID : EV_GENERIC, PAIR>,
                                      cudaFuncAttributeMaxDynamicSharedMemorySize, bytes));
  F5B_CHECK_CUDA(cudaFuncSetAttribute(tc_gemm_kernel<BN, STAGES, SPECIAL ? EV_GELU16 : EV_GENERIC, PAIR>,
                                      cudaFuncAttributeMaxDynamicSharedMemorySize, bytes));
}
void tc_gemm_init_kernels() {
  config_one<64, 6, 0>();
  config_one<128, 5, 0>();
  config_one<192, 4, 0>();
  config_one<256, 3, 0>();
  config_one<128, 6, 1>();
  config_one<192, 5, 1>();
  config_one<256, 4, 1>();
}

// Tile width, CTA pairing and split-K factor.  Cost model from the in-situ timeline (tools/ktrace.py, B200): the mainloop
// is paced by the bytes an SM takes in per 64-deep k-block (A 16 KB + B 128 B per tile column, ~64 B/clk): ~0.26 / 0.28 /
// 0.35 / 0.43 us for 64 / 128 / 192 / 256-wide tiles; a CTA pair halves the B bytes per CTA (0.20 / 0.23 / 0.27 us, the
// last one at the tensor pipe's own floor).  A CTA runs its units back to back with the epilogue of all but the last
// hidden behind the next unit's MMAs; launch, pipeline fill and store drain do not depend on the tiling.
static void pick_tiling(long long tiles_m_per_batch, int batches, int N, int K, int sms, bool can_split, bool can_pair,
                        int force_bn, int force_pair, int force_splits, int* bn_out, int* pair_out, int* splits_out) {
  const int cand[4] = {64, 128, 192, 256};
  // pair row: ratios to the single-CTA row measured on B200 after the elect.sync issue fix (profiles/r2_gemm_sweep.log,
  // M = 32768: 0.99 / 0.92 / 0.88 per k-block round at 128 / 192 / 256 columns, each round covering 256 rows)
  const double kb_us[2][4] = {{0.26, 0.28, 0.35, 0.43}, {9.9, 0.277, 0.322, 0.378}};
  const double epi_us[4] = {0.8, 1.4, 2.4, 2.5};
  const int num_kb = (K + BK - 1) / BK;
  double best_cost = 1e30;
  *bn_out = 128;
  *pair_out = 0;
  *splits_out = 1;
  for (int pair = 0; pair <= 1; ++pair) {
    if (pair && !can_pair) continue;
    if (force_pair >= 0 && pair != force_pair) continue;
    for (int i = pair ? 1 : 0; i < 4; ++i) {
      if (force_bn && cand[i] != force_bn) continue;
      for (int sp = 1; sp <= 4; sp *= 2) {
        if (force_splits && sp != force_splits) continue;
        if (sp > 1 && (pair || !can_split || num_kb / sp < 4)) continue;
        const long long mrows = pair ? (tiles_m_per_batch + 1) / 2 : tiles_m_per_batch;
        const long long units = mrows * batches * ((N + cand[i] - 1) / cand[i]) * sp;
        const long long workers = pair ? sms / 2 : sms;
        const long long rounds = (units + workers - 1) / workers;
        const double cost = rounds * (kb_us[pair][i] * ((num_kb + sp - 1) / sp) + 0.15) + epi_us[i] + (pair ? 0.6 : 0.0);
        if (cost < best_cost) {
          best_cost = cost;
          *bn_out = cand[i];
          *pair_out = pair;
          *splits_out = sp;
        }
      }
    }
  }
}

void tc_gemm(cudaStream_t st, int sm_count, const bf16* A, long long lda, int rows_per_batch, int batches,
             const bf16* W, long long ldw, int N, int K, const EpiParams& ep, const TcConv* conv, int force_bn,
             int force_splits, int fp16_operands) {
  F5B_REQUIRE(rows_per_batch > 0 && batches > 0 && N > 0 && K > 0, "empty GEMM");
  TcGemmShape s{};
  s.rows_per_batch = rows_per_batch;
  s.batches = batches;
  s.tiles_m_per_batch = (rows_per_batch + BM - 1) / BM;
  s.N = N;
  int bn;
  int pair = 0;  // CTA pairs (cta_group::2)
  int a_cols;  // width of the A matrix (TMA dim 0)
  if (conv != nullptr) {
    F5B_REQUIRE(conv->cg % BK == 0, "tensor-core conv needs channels-per-group % 64 == 0");
    F5B_REQUIRE(K == conv->taps * conv->cg, "conv K mismatch");
    bn = 64;
    s.splits = 1;
    s.conv = 1;
    s.conv_kb_per_tap = conv->cg / BK;
    s.conv_pad = conv->pad;
    s.conv_cg = conv->cg;
    a_cols = conv->channels;
  } else {
    // split-K meets in an L2 reduce-add whose order is not fixed: results would differ in the last ulp from run to run
    // (and flip 16-bit roundings downstream), so it is only used on request (F5B_SPLIT_K=1 / force_splits)
    static const bool allow_split = std::getenv("F5B_SPLIT_K") != nullptr;
    const bool can_split = ep.kind == EPI_RESID && ep.act == ACT_NONE && (allow_split || force_splits > 1);
    // force_bn: 0 = model's choice, 64..256 = that width on single CTAs, 1000 + width = that width on CTA pairs
    // CTA pairs win from M ~ 5 k rows on (qkv at M = 32768: 1365 vs 1200 TFLOP/s) since the single-lane roles issue
    // through elect.sync; F5B_NO_PAIR=1 switches them off
    const int fpair = force_bn >= 1000 ? 1 : force_bn > 0 ? 0 : -1;
    static const bool no_pair = std::getenv("F5B_NO_PAIR") != nullptr;
    pick_tiling(s.tiles_m_per_batch, batches, N, K, sm_count, can_split, !no_pair || fpair == 1, force_bn % 1000, fpair, force_splits,
                &bn, &pair, &s.splits);
    a_cols = K;
  }
  s.tiles_n = (N + bn - 1) / bn;
  s.num_kb = (K + BK - 1) / BK;
  const int fp16 = fp16_operands || (conv != nullptr && conv->fp16);
  s.idesc = fp16 ? umma_idesc_f16(BM * (1 + pair), bn, 0, 0) : umma_idesc_bf16(BM * (1 + pair), bn, 0, 0);

  CUtensorMap ta, tb;
  {
    uint64_t dims[3] = {(uint64_t)a_cols, (uint64_t)rows_per_batch, (uint64_t)batches};
    uint64_t strides[2] = {(uint64_t)lda * 2, (uint64_t)lda * 2 * (uint64_t)rows_per_batch};
    uint32_t box[3] = {BK, BM, 1};
    make_tmap_bf16(&ta, A, 3, dims, strides, box, fp16);
  }
  {
    uint64_t dims[2] = {(uint64_t)K, (uint64_t)N};
    uint64_t strides[1] = {(uint64_t)ldw * 2};
    uint32_t box[2] = {BK, (uint32_t)(pair ? bn / 2 : bn)};
    make_tmap_bf16(&tb, W, 2, dims, strides, box, fp16);
  }
  // output map: 3-D like A so that tile rows beyond a batch element (and the rows dropped by row_skip, which land at
  // coordinate -1) are clipped by the TMA engine; box = 128 rows x 128 bytes, 128B-swizzled like the staging buffer
  CUtensorMap to;
  {
    F5B_REQUIRE(ep.addend == nullptr, "tensor-core epilogue has no addend (use the residual form)");
    const bool resid = ep.kind == EPI_RESID;
    const int et = resid ? OUT_F32 : ep.out_type;
    const int es = et == OUT_F32 ? 4 : 2;
    const void* base = resid ? (const void*)ep.resid : (const void*)ep.out;
    const long long ld = resid ? ep.ldr : ep.ldo;
    const int ncols = ep.n_valid > 0 && ep.n_valid < N ? ep.n_valid : N;
    F5B_REQUIRE(base != nullptr && ld > 0, "GEMM output missing");
    F5B_REQUIRE(ep.row_skip == 0 || ep.rows_per_batch == rows_per_batch, "row_skip needs per-batch tiling");
    const int out_rows = rows_per_batch - ep.row_skip;
    uint64_t dims[3] = {(uint64_t)ncols, (uint64_t)out_rows, (uint64_t)batches};
    uint64_t strides[2] = {(uint64_t)ld * es, (uint64_t)ld * es * (uint64_t)out_rows};
    uint32_t box[3] = {(uint32_t)(128 / es), BM, 1};
    make_tmap(&to, base, 3, dims, strides, box, et);
  }
  if (sync_debug())
    fprintf(stderr, "tc_gemm rows %d x %d N %d K %d bn %d pair %d splits %d conv %d kind %d out_type %d act %d row_skip %d\n",
            rows_per_batch, batches, N, K, bn, pair, s.splits, s.conv, ep.kind, ep.out_type, ep.act, ep.row_skip);
  if (pair) {
    const long long units = (long long)((s.tiles_m_per_batch + 1) / 2) * batches * s.tiles_n;
    const int pairs = (int)(units < sm_count / 2 ? units : sm_count / 2);
    if (bn == 128) launch_cfg<128, 6, 1>(st, 2 * pairs, ta, tb, to, s, ep);
    else if (bn == 192) launch_cfg<192, 5, 1>(st, 2 * pairs, ta, tb, to, s, ep);
    else launch_cfg<256, 4, 1>(st, 2 * pairs, ta, tb, to, s, ep);
    return;
  }
  long long total = (long long)s.tiles_m_per_batch * batches * s.tiles_n * s.splits;
  int grid = (int)(total < sm_count ? total : sm_count);
  if (bn == 64)
    launch_cfg<64, 6, 0>(st, grid, ta, tb, to, s, ep);
  else if (bn == 128)
    launch_cfg<128, 5, 0>(st, grid, ta, tb, to, s, ep);
  else if (bn == 192)
    launch_cfg<192, 4, 0>(st, grid, ta, tb, to, s, ep);
  else
    launch_cfg<256, 3, 0>(st, grid, ta, tb, to, s, ep);
}

}  // namespace f5b
