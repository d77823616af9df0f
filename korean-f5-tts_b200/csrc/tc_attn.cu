// Non-causal multi-head attention, head_dim 64, on tcgen05 tensor cores with TMEM accumulators.
//
// Replaces AttnProcessor.__call__'s F.scaled_dot_product_attention (reference src/f5_tts/model/modules.py:499-508;
// flash-attn variant :510-531) for the DiT / UNetT blocks.  q, k (already rotated by the QKV GEMM epilogue) and v
// live in one bf16 [batches*rows, 3*heads*64] buffer; one CTA owns one (128-query tile, head, batch element).
//
//   warps 0..3 : softmax -- thread t owns query row t (= TMEM lane t): tcgen05.ld of S, online max / exp2 / sum in
//                fp32 registers, P written to smem as bf16 in the 128B-swizzled K-major UMMA layout, O kept in
//                registers and updated from the per-block P*V product read back from TMEM
//   warp 4     : TMA producer (Q once; K_j, V_j double buffered)
//   warp 5     : MMA issuer: S = Q K_j^T (4 x 128x128x16), O_j = P V_j (8 x 128x64x16, V read MN-major)
// Two CTAs fit per SM (112 KB smem, 256 TMEM columns each) so one CTA's softmax overlaps the other's MMAs.
#include "common.cuh"
#include "sm100_ptx.cuh"
#include "tc_gemm.h"

namespace f5b {

using namespace ptx;

static constexpr int AT_BQ = 128, AT_BKV = 128, AT_D = 64;
static constexpr int AT_TILE_BYTES = 128 * 64 * 2;  // 16 KB: Q, K_j, V_j, and each of the two P panels
static constexpr int AT_SMEM = AT_TILE_BYTES * (1 + 2 + 2 + 2) + 1024 + 128;
static constexpr uint32_t AT_TMEM_COLS = 256;  // S: [0,128), O_j: [128,192)

__global__ void __launch_bounds__(192, 2)
tc_attn_kernel(const __grid_constant__ CUtensorMap tm, bf16* __restrict__ out, int rows, int heads,
               const int* __restrict__ kv_len, float scale_log2) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
  uint8_t* sQ = smem;
  uint8_t* sK = sQ + AT_TILE_BYTES;       // 2 stages
  uint8_t* sV = sK + 2 * AT_TILE_BYTES;   // 2 stages
  uint8_t* sP = sV + 2 * AT_TILE_BYTES;   // 2 panels of 64 kv columns
  uint64_t* bars = reinterpret_cast<uint64_t*>(sP + 2 * AT_TILE_BYTES);
  uint64_t* q_full = bars;
  uint64_t* kv_full = bars + 1;   // [2]
  uint64_t* kv_empty = bars + 3;  // [2]
  uint64_t* s_full = bars + 5;
  uint64_t* p_full = bars + 6;
  uint64_t* o_full = bars + 7;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 8);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int q0 = blockIdx.x * AT_BQ, h = blockIdx.y, b = blockIdx.z;
  const int inner = heads * AT_D;
  const int L = kv_len ? kv_len[b] : rows;
  const int nkv = (L + AT_BKV - 1) / AT_BKV;

  if (threadIdx.x == 0) {
    prefetch_tmap(&tm);
    mbar_init(q_full, 1);
    for (int i = 0; i < 2; ++i) {
      mbar_init(&kv_full[i], 1);
      mbar_init(&kv_empty[i], 1);
    }
    mbar_init(s_full, 1);
    mbar_init(p_full, 128);
    mbar_init(o_full, 1);
    fence_barrier_init();
  }
  if (warp == 5) tmem_alloc(tmem_slot, AT_TMEM_COLS);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  const uint32_t tS = tmem_base, tO = tmem_base + 128;

  if (warp == 4 && lane == 0) {
    // ------------------------------- TMA producer -------------------------------
    mbar_expect_tx(q_full, AT_TILE_BYTES);
    tma_load_3d(sQ, &tm, q_full, h * AT_D, q0, b);
    for (int j = 0; j < nkv; ++j) {
      const int st = j & 1;
      mbar_wait(&kv_empty[st], ((j >> 1) & 1) ^ 1);
      mbar_expect_tx(&kv_full[st], 2 * AT_TILE_BYTES);
      tma_load_3d(sK + st * AT_TILE_BYTES, &tm, &kv_full[st], inner + h * AT_D, j * AT_BKV, b);
      tma_load_3d(sV + st * AT_TILE_BYTES, &tm, &kv_full[st], 2 * inner + h * AT_D, j * AT_BKV, b);
    }
  } else if (warp == 5 && lane == 0) {
    // ------------------------------- MMA issuer ---------------------------------
    constexpr uint32_t IDESC_S = umma_idesc_bf16(128, 128, 0, 0);
    constexpr uint32_t IDESC_O = umma_idesc_bf16(128, 64, 0, 1);  // B = V is MN-major
    mbar_wait(q_full, 0);
    const uint32_t q_addr = smem_u32(sQ), p_addr = smem_u32(sP);
    for (int j = 0; j < nkv; ++j) {
      const int st = j & 1;
      mbar_wait(&kv_full[st], (j >> 1) & 1);
      tc_fence_after();
      const uint32_t k_addr = smem_u32(sK + st * AT_TILE_BYTES);
      const uint32_t v_addr = smem_u32(sV + st * AT_TILE_BYTES);
#pragma unroll
      for (int k = 0; k < AT_D / 16; ++k)
        umma_ss(tS, umma_desc_sw128(q_addr + k * 32, 16, 1024), umma_desc_sw128(k_addr + k * 32, 16, 1024), IDESC_S,
                k != 0 ? 1u : 0u);
      umma_commit(s_full);
      mbar_wait(p_full, j & 1);  // P_j is in smem, S and O_{j-1} have been read out of TMEM
      tc_fence_after();
#pragma unroll
      for (int k = 0; k < AT_BKV / 16; ++k)
        umma_ss(tO, umma_desc_sw128(p_addr + (k >> 2) * AT_TILE_BYTES + (k & 3) * 32, 16, 1024),
                umma_desc_sw128(v_addr + k * 16 * 128, 1024, 1024), IDESC_O, k != 0 ? 1u : 0u);
      umma_commit(o_full);
      umma_commit(&kv_empty[st]);
    }
  } else if (warp < 4) {
    // ------------------------------- softmax / output ---------------------------
    const int row = warp * 32 + lane;
    const uint32_t lane_off = static_cast<uint32_t>(warp * 32) << 16;
    float o[AT_D];
#pragma unroll
    for (int d = 0; d < AT_D; ++d) o[d] = 0.0f;
    float mx = -INFINITY, l = 0.0f;
    uint8_t* p_row = sP + row * 128;
    const int sw = row & 7;

    for (int j = 0; j < nkv; ++j) {
      mbar_wait(s_full, j & 1);
      tc_fence_after();
      const int kv_base = j * AT_BKV;
      const bool tail = kv_base + AT_BKV > L;
      // pass 1: row max
      float cmax = -INFINITY;
#pragma unroll 1
      for (int c = 0; c < AT_BKV; c += 32) {
        float v[32];
        tmem_ld_32x32(tS + lane_off + c, v);
        tmem_ld_wait();
#pragma unroll
        for (int i = 0; i < 32; ++i) {
          float sv = (tail && kv_base + c + i >= L) ? -INFINITY : v[i];
          cmax = fmaxf(cmax, sv);
        }
      }
      const float mnew = fmaxf(mx, cmax);
      const float alpha = exp2f((mx - mnew) * scale_log2);
      const float moff = mnew * scale_log2;
      float psum = 0.0f;
      // pass 2: p = exp2(s*scale - m*scale) -> bf16 -> swizzled smem
#pragma unroll 1
      for (int c = 0; c < AT_BKV; c += 32) {
        float v[32];
        tmem_ld_32x32(tS + lane_off + c, v);
        tmem_ld_wait();
#pragma unroll
        for (int i = 0; i < 32; ++i) {
          float p = exp2f(fmaf(v[i], scale_log2, -moff));
          if (tail && kv_base + c + i >= L) p = 0.0f;
          psum += p;
          v[i] = p;
        }
        uint8_t* panel = p_row + (c >> 6) * AT_TILE_BYTES;
        const int ch0 = (c & 63) >> 3;
#pragma unroll
        for (int g = 0; g < 4; ++g) {
          uint4 pk;
          pk.x = pack_bf16x2(v[g * 8 + 0], v[g * 8 + 1]);
          pk.y = pack_bf16x2(v[g * 8 + 2], v[g * 8 + 3]);
          pk.z = pack_bf16x2(v[g * 8 + 4], v[g * 8 + 5]);
          pk.w = pack_bf16x2(v[g * 8 + 6], v[g * 8 + 7]);
          *reinterpret_cast<uint4*>(panel + (((ch0 + g) ^ sw) << 4)) = pk;
        }
      }
      l = l * alpha + psum;
      mx = mnew;
#pragma unroll
      for (int d = 0; d < AT_D; ++d) o[d] *= alpha;
      fence_proxy_async_smem();  // generic-proxy smem writes -> visible to the tensor core (async proxy)
      tc_fence_before();
      mbar_arrive(p_full);
      // O += P_j V_j
      mbar_wait(o_full, j & 1);
      tc_fence_after();
#pragma unroll
      for (int c = 0; c < AT_D; c += 32) {
        float v[32];
        tmem_ld_32x32(tO + lane_off + c, v);
        tmem_ld_wait();
#pragma unroll
        for (int i = 0; i < 32; ++i) o[c + i] += v[i];
      }
    }
    const int r = q0 + row;
    if (r < rows) {
      const float inv = 1.0f / l;
      bf16* op = out + ((long long)b * rows + r) * inner + h * AT_D;
#pragma unroll
      for (int d = 0; d < AT_D; d += 8) {
        uint4 pk;
        pk.x = pack_bf16x2(o[d] * inv, o[d + 1] * inv);
        pk.y = pack_bf16x2(o[d + 2] * inv, o[d + 3] * inv);
        pk.z = pack_bf16x2(o[d + 4] * inv, o[d + 5] * inv);
        pk.w = pack_bf16x2(o[d + 6] * inv, o[d + 7] * inv);
        *reinterpret_cast<uint4*>(op + d) = pk;
      }
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 5) {
    tc_fence_after();
    tmem_dealloc(tmem_base, AT_TMEM_COLS);
  }
}

void tc_attention(cudaStream_t st, const bf16* qkv, bf16* out, int batches, int rows, int heads, const int* kv_len,
                  float scale) {
  const int inner = heads * AT_D;
  CUtensorMap tm;
  uint64_t dims[3] = {(uint64_t)3 * inner, (uint64_t)rows, (uint64_t)batches};
  uint64_t strides[2] = {(uint64_t)3 * inner * 2, (uint64_t)3 * inner * 2 * (uint64_t)rows};
  uint32_t box[3] = {AT_D, 128, 1};
  make_tmap_bf16(&tm, qkv, 3, dims, strides, box);
  dim3 grid((rows + AT_BQ - 1) / AT_BQ, heads, batches);
  tc_attn_kernel<<<grid, 192, AT_SMEM, st>>>(tm, out, rows, heads, kv_len, scale * 1.4426950408889634f);
  F5B_POST_LAUNCH();
}

void tc_init_kernels() {
  F5B_CHECK_CUDA(cudaFuncSetAttribute(tc_attn_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, AT_SMEM));
}

}  // namespace f5b
