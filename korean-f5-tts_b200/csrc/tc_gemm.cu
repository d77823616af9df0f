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
//   warps 2..5    : epilogue       -- tcgen05.ld 32x32b, fused bias / RoPE / activation / gate+residual,
//                                     direct global stores; overlaps the next tile's MMAs
#include "common.cuh"
#include "sm100_ptx.cuh"
#include "tc_gemm.h"

#include <cudaTypedefs.h>
#include <mutex>

namespace f5b {

using namespace ptx;

static constexpr int BM = 128;
static constexpr int BK = 64;  // 64 bf16 = 128 B = one swizzle row
static constexpr int A_STAGE_BYTES = BM * BK * 2;

template <int BN, int STAGES>
__global__ void __launch_bounds__(192, 1)
tc_gemm_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
               const TcGemmShape s, const EpiParams ep) {
  constexpr int B_STAGE_BYTES = BN * BK * 2;
  constexpr uint32_t STAGE_TX = A_STAGE_BYTES + B_STAGE_BYTES;
  constexpr uint32_t ACC_COLS = 2 * BN;

  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
  uint8_t* sA = smem;
  uint8_t* sB = smem + STAGES * A_STAGE_BYTES;
  uint64_t* full = reinterpret_cast<uint64_t*>(smem + STAGES * (A_STAGE_BYTES + B_STAGE_BYTES));
  uint64_t* empty = full + STAGES;
  uint64_t* tfull = empty + STAGES;
  uint64_t* tempty = tfull + 2;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tempty + 2);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const int tiles_m = s.tiles_m_per_batch * s.batches;
  const int total_tiles = tiles_m * s.tiles_n;

  if (threadIdx.x == 0) {
    prefetch_tmap(&tmA);
    prefetch_tmap(&tmB);
    for (int i = 0; i < STAGES; ++i) {
      mbar_init(&full[i], 1);
      mbar_init(&empty[i], 1);
    }
    for (int i = 0; i < 2; ++i) {
      mbar_init(&tfull[i], 1);
      mbar_init(&tempty[i], 4);
    }
    fence_barrier_init();
  }
  if (warp == 1) tmem_alloc(tmem_slot, ACC_COLS);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0 && lane == 0) {
    // ------------------------------- TMA producer -------------------------------
    uint32_t it = 0;
    for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x) {
      const int tm = tile / s.tiles_n, tn = tile - tm * s.tiles_n;
      const int bz = tm / s.tiles_m_per_batch;
      const int r0 = (tm - bz * s.tiles_m_per_batch) * BM;
      const int n0 = tn * BN;
      for (int kb = 0; kb < s.num_kb; ++kb, ++it) {
        const uint32_t st = it % STAGES, ph = (it / STAGES) & 1u;
        mbar_wait(&empty[st], ph ^ 1u);
        mbar_expect_tx(&full[st], STAGE_TX);
        int a_c0 = kb * BK, a_c1 = r0;
        if (s.conv) {
          const int tap = kb / s.conv_kb_per_tap, sub = kb - tap * s.conv_kb_per_tap;
          a_c0 = (n0 / s.conv_cg) * s.conv_cg + sub * BK;
          a_c1 = r0 + tap - s.conv_pad;
        }
        tma_load_3d(sA + st * A_STAGE_BYTES, &tmA, &full[st], a_c0, a_c1, bz);
        tma_load_2d(sB + st * B_STAGE_BYTES, &tmB, &full[st], kb * BK, n0);
      }
    }
  } else if (warp == 1 && lane == 0) {
    // ------------------------------- MMA issuer ---------------------------------
    uint32_t it = 0, tcount = 0;
    for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x, ++tcount) {
      const uint32_t acc = tcount & 1u, aph = (tcount >> 1) & 1u;
      mbar_wait(&tempty[acc], aph ^ 1u);
      tc_fence_after();
      const uint32_t d_tmem = tmem_base + acc * BN;
      for (int kb = 0; kb < s.num_kb; ++kb, ++it) {
        const uint32_t st = it % STAGES, ph = (it / STAGES) & 1u;
        mbar_wait(&full[st], ph);
        tc_fence_after();
        const uint32_t a_addr = smem_u32(sA + st * A_STAGE_BYTES);
        const uint32_t b_addr = smem_u32(sB + st * B_STAGE_BYTES);
#pragma unroll
        for (int k = 0; k < BK / 16; ++k) {
          const uint64_t ad = umma_desc_sw128(a_addr + k * 32, 16, 1024);
          const uint64_t bd = umma_desc_sw128(b_addr + k * 32, 16, 1024);
          umma_ss(d_tmem, ad, bd, s.idesc, (kb | k) != 0 ? 1u : 0u);
        }
        umma_commit(&empty[st]);  // frees the smem slot once these MMAs have read it
      }
      umma_commit(&tfull[acc]);  // accumulator complete
    }
  } else if (warp >= 2) {
    // ------------------------------- epilogue -----------------------------------
    const int quarter = warp & 3;  // TMEM lanes [32*quarter, 32*quarter+32) are this warp's
    uint32_t tcount = 0;
    for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x, ++tcount) {
      const int tm = tile / s.tiles_n, tn = tile - tm * s.tiles_n;
      const int bz = tm / s.tiles_m_per_batch;
      const int r0 = (tm - bz * s.tiles_m_per_batch) * BM;
      const int n0 = tn * BN;
      const uint32_t acc = tcount & 1u, aph = (tcount >> 1) & 1u;
      mbar_wait(&tfull[acc], aph);
      tc_fence_after();
      const int r = r0 + quarter * 32 + lane;
      const bool row_ok = r < s.rows_per_batch;
      const long long m = (long long)bz * s.rows_per_batch + r;
      const uint32_t t_row = tmem_base + acc * BN + (static_cast<uint32_t>(quarter * 32) << 16);
#pragma unroll 1
      for (int c = 0; c < BN; c += 32) {
        float v[32];
        tmem_ld_32x32(t_row + c, v);
        tmem_ld_wait();
        if (row_ok && n0 + c < s.N) epi_apply<32>(ep, m, n0 + c, v);
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&tempty[acc]);
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc(tmem_base, ACC_COLS);
  }
}

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
  CUresult r = g_encode(out, fp16 ? CU_TENSOR_MAP_DATA_TYPE_FLOAT16 : CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, (cuuint32_t)rank, const_cast<void*>(base), gd, gs, bx,
                        es, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                        CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    set_error("cuTensorMapEncodeTiled failed with CUresult " + std::to_string((int)r));
    throw Err{F5B_ERR_CUDA};
  }
}

template <int BN, int STAGES>
static void launch_cfg(cudaStream_t st, int grid, const CUtensorMap& ta, const CUtensorMap& tb, const TcGemmShape& s,
                       const EpiParams& ep) {
  constexpr int smem = STAGES * (A_STAGE_BYTES + BN * BK * 2) + 1024 + 256;
  tc_gemm_kernel<BN, STAGES><<<grid, 192, smem, st>>>(ta, tb, s, ep);
  F5B_POST_LAUNCH();
}

template <int BN, int STAGES>
static void config_one() {
  constexpr int smem = STAGES * (A_STAGE_BYTES + BN * BK * 2) + 1024 + 256;
  F5B_CHECK_CUDA(cudaFuncSetAttribute(tc_gemm_kernel<BN, STAGES>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
}
void tc_gemm_init_kernels() {
  config_one<64, 8>();
  config_one<128, 6>();
  config_one<256, 4>();
}

static int pick_bn(long long tiles_m, int N, int K, int sms, int force) {
  if (force == 64 || force == 128 || force == 256) return force;
  int best = 128;
  double best_cost = 1e30;
  const int cand[3] = {64, 128, 256};
  for (int i = 0; i < 3; ++i) {
    int bn = cand[i];
    long long tiles = tiles_m * ((N + bn - 1) / bn);
    long long waves = (tiles + sms - 1) / sms;
    double per_k16 = bn / 2.0 > 32.0 + bn / 4.0 ? bn / 2.0 : 32.0 + bn / 4.0;  // MMA issue vs smem operand reads
    double tile_cost = per_k16 * (K / 16.0) + 1200.0 * bn / 128.0 / (waves > 1 ? 4.0 : 1.0);
    double cost = waves * tile_cost;
    if (cost < best_cost) {
      best_cost = cost;
      best = bn;
    }
  }
  return best;
}

void tc_gemm(cudaStream_t st, int sm_count, const bf16* A, long long lda, int rows_per_batch, int batches,
             const bf16* W, long long ldw, int N, int K, const EpiParams& ep, const TcConv* conv, int force_bn) {
  F5B_REQUIRE(rows_per_batch > 0 && batches > 0 && N > 0 && K > 0, "empty GEMM");
  TcGemmShape s{};
  s.rows_per_batch = rows_per_batch;
  s.batches = batches;
  s.tiles_m_per_batch = (rows_per_batch + BM - 1) / BM;
  s.N = N;
  int bn;
  int a_cols;  // width of the A matrix (TMA dim 0)
  if (conv != nullptr) {
    F5B_REQUIRE(conv->cg % BK == 0, "tensor-core conv needs channels-per-group % 64 == 0");
    F5B_REQUIRE(K == conv->taps * conv->cg, "conv K mismatch");
    bn = 64;
    s.conv = 1;
    s.conv_kb_per_tap = conv->cg / BK;
    s.conv_pad = conv->pad;
    s.conv_cg = conv->cg;
    a_cols = conv->channels;
  } else {
    bn = pick_bn((long long)s.tiles_m_per_batch * batches, N, K, sm_count, force_bn);
    a_cols = K;
  }
  s.tiles_n = (N + bn - 1) / bn;
  s.num_kb = (K + BK - 1) / BK;
  const int fp16 = conv != nullptr && conv->fp16;
  s.idesc = fp16 ? umma_idesc_f16(BM, bn, 0, 0) : umma_idesc_bf16(BM, bn, 0, 0);

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
    uint32_t box[2] = {BK, (uint32_t)bn};
    make_tmap_bf16(&tb, W, 2, dims, strides, box, fp16);
  }
  long long total = (long long)s.tiles_m_per_batch * batches * s.tiles_n;
  int grid = (int)(total < sm_count ? total : sm_count);
  if (bn == 64)
    launch_cfg<64, 8>(st, grid, ta, tb, s, ep);
  else if (bn == 128)
    launch_cfg<128, 6>(st, grid, ta, tb, s, ep);
  else
    launch_cfg<256, 4>(st, grid, ta, tb, s, ep);
}

}  // namespace f5b
