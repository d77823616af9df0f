// fp32 CUDA-core contraction kernels: the "fp32 mode" (correctness mode, parity 1e-3 against the reference's fp32
// CPU path) of the GEMM / grouped conv / attention, plus the skinny GEMM used for the per-sample AdaLN tables.
// Same fused epilogue as the tensor-core kernels (common.cuh).
#include "common.cuh"
#include "tc_gemm.h"

namespace f5b {

// ---------------------------------------------------------------------------------------------
// C[M,N] = A[M,K] * W[N,K]^T, 128x128x16 tiles, 256 threads, 8x8 per thread (two 4-wide column groups).
// ---------------------------------------------------------------------------------------------
static constexpr int SBM = 128, SBN = 128, SBK = 16, SPAD = 4;

__global__ void __launch_bounds__(256)
simt_gemm_kernel(const float* __restrict__ A, long long lda, long long M, const float* __restrict__ W, long long ldw,
                 int N, int K, const EpiParams ep) {
  __shared__ float As[SBK][SBM + SPAD];
  __shared__ float Bs[SBK][SBN + SPAD];
  pdl_trigger();
  pdl_wait();
  const int tid = threadIdx.x;
  const int tx = tid & 15, ty = tid >> 4;
  const long long m0 = (long long)blockIdx.y * SBM;
  const int n0 = blockIdx.x * SBN;

  float acc[8][8];
#pragma unroll
  for (int i = 0; i < 8; ++i)
#pragma unroll
    for (int j = 0; j < 8; ++j) acc[i][j] = 0.0f;

  // each thread stages 2 float4 of A and 2 of W per k-block
  const int lrow = tid >> 2;        // 0..63
  const int lk = (tid & 3) * 4;     // 0,4,8,12
  float4 ra[2], rb[2];

  auto load_regs = [&](int k0) {
#pragma unroll
    for (int i = 0; i < 2; ++i) {
      long long r = m0 + lrow + i * 64;
      int k = k0 + lk;
      ra[i] = (r < M && k < K) ? *reinterpret_cast<const float4*>(A + r * lda + k) : make_float4(0, 0, 0, 0);
      int n = n0 + lrow + i * 64;
      rb[i] = (n < N && k < K) ? *reinterpret_cast<const float4*>(W + (long long)n * ldw + k) : make_float4(0, 0, 0, 0);
    }
  };
  auto store_smem = [&]() {
#pragma unroll
    for (int i = 0; i < 2; ++i) {
      int r = lrow + i * 64;
      As[lk + 0][r] = ra[i].x; As[lk + 1][r] = ra[i].y; As[lk + 2][r] = ra[i].z; As[lk + 3][r] = ra[i].w;
      Bs[lk + 0][r] = rb[i].x; Bs[lk + 1][r] = rb[i].y; Bs[lk + 2][r] = rb[i].z; Bs[lk + 3][r] = rb[i].w;
    }
  };

  load_regs(0);
  for (int k0 = 0; k0 < K; k0 += SBK) {
    store_smem();
    __syncthreads();
    if (k0 + SBK < K) load_regs(k0 + SBK);
#pragma unroll
    for (int k = 0; k < SBK; ++k) {
      float a[8], b[8];
      float4 a0 = *reinterpret_cast<const float4*>(&As[k][ty * 4]);
      float4 a1 = *reinterpret_cast<const float4*>(&As[k][ty * 4 + 64]);
      float4 b0 = *reinterpret_cast<const float4*>(&Bs[k][tx * 4]);
      float4 b1 = *reinterpret_cast<const float4*>(&Bs[k][tx * 4 + 64]);
      a[0] = a0.x; a[1] = a0.y; a[2] = a0.z; a[3] = a0.w; a[4] = a1.x; a[5] = a1.y; a[6] = a1.z; a[7] = a1.w;
      b[0] = b0.x; b[1] = b0.y; b[2] = b0.z; b[3] = b0.w; b[4] = b1.x; b[5] = b1.y; b[6] = b1.z; b[7] = b1.w;
#pragma unroll
      for (int i = 0; i < 8; ++i)
#pragma unroll
        for (int j = 0; j < 8; ++j) acc[i][j] = fmaf(a[i], b[j], acc[i][j]);
    }
    __syncthreads();
  }

#pragma unroll
  for (int i = 0; i < 8; ++i) {
    long long m = m0 + ty * 4 + (i & 3) + (i >> 2) * 64;
    if (m >= M) continue;
#pragma unroll
    for (int h = 0; h < 2; ++h) {
      int n = n0 + tx * 4 + h * 64;
      if (n < N) epi_apply<4>(ep, m, n, &acc[i][h * 4]);
    }
  }
}

// ---------------------------------------------------------------------------------------------
// Skinny GEMM, M <= 16 rows per pass: one warp per 4 output columns, A chunk resident in smem.
// ---------------------------------------------------------------------------------------------
static constexpr int SK_KC = 1024;  // K chunk held in smem: 16 x 1024 fp32 = 64 KB
static constexpr int SK_SMEM = (16 * SK_KC + 8 * 16 * 4) * (int)sizeof(float);

__global__ void __launch_bounds__(256)
skinny_gemm_kernel(const float* __restrict__ A, long long lda, int M, const float* __restrict__ W, long long ldw,
                   int N, int K, const EpiParams ep, long long m_base) {
  extern __shared__ float sk_smem[];
  float* As = sk_smem;                       // [16][SK_KC]
  float* red = sk_smem + 16 * SK_KC;         // [8 warps][16][4]
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int n4 = (blockIdx.x * 8 + warp) * 4;
  float acc[16][4];
#pragma unroll
  for (int i = 0; i < 16; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) acc[i][j] = 0.0f;

  for (int kc = 0; kc < K; kc += SK_KC) {
    const int kw = min(SK_KC, K - kc);
    __syncthreads();
    for (int idx = threadIdx.x; idx < 16 * (SK_KC / 4); idx += 256) {
      int r = idx / (SK_KC / 4), k = (idx % (SK_KC / 4)) * 4;
      float4 v = make_float4(0, 0, 0, 0);
      if (r < M && k < kw) v = *reinterpret_cast<const float4*>(A + (m_base + r) * lda + kc + k);
      *reinterpret_cast<float4*>(As + r * SK_KC + k) = v;
    }
    __syncthreads();
    if (n4 < N) {
      for (int k = lane * 4; k < kw; k += 128) {
        float4 w[4];
#pragma unroll
        for (int j = 0; j < 4; ++j)
          w[j] = (n4 + j < N) ? __ldg(reinterpret_cast<const float4*>(W + (long long)(n4 + j) * ldw + kc + k))
                              : make_float4(0, 0, 0, 0);
#pragma unroll
        for (int i = 0; i < 16; ++i) {
          float4 a = *reinterpret_cast<const float4*>(As + i * SK_KC + k);
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            acc[i][j] = fmaf(a.x, w[j].x, acc[i][j]);
            acc[i][j] = fmaf(a.y, w[j].y, acc[i][j]);
            acc[i][j] = fmaf(a.z, w[j].z, acc[i][j]);
            acc[i][j] = fmaf(a.w, w[j].w, acc[i][j]);
          }
        }
      }
    }
  }
#pragma unroll
  for (int i = 0; i < 16; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      float v = warp_sum(acc[i][j]);
      if (lane == 0) red[(warp * 16 + i) * 4 + j] = v;
    }
  __syncwarp();
  if (n4 < N && lane < M) {
    float v[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) v[j] = red[(warp * 16 + lane) * 4 + j];
    epi_apply<4>(ep, m_base + lane, n4, v);
  }
}

void simt_gemm(cudaStream_t st, const float* A, long long lda, long long M, const float* W, long long ldw, int N,
               int K, const EpiParams& ep) {
  F5B_REQUIRE(M > 0 && N > 0 && K > 0, "empty GEMM");
  F5B_REQUIRE(K % 4 == 0 && lda % 4 == 0 && ldw % 4 == 0, "fp32 GEMM needs K, lda, ldw multiples of 4");
  F5B_REQUIRE(((uintptr_t)A & 15) == 0 && ((uintptr_t)W & 15) == 0, "fp32 GEMM operands must be 16-byte aligned");
  if (M <= 32) {
    int blocks = (N + 31) / 32;
    for (long long mb = 0; mb < M; mb += 16) {
      int mm = (int)(M - mb < 16 ? M - mb : 16);
      skinny_gemm_kernel<<<blocks, 256, SK_SMEM, st>>>(A, lda, mm, W, ldw, N, K, ep, mb);
      F5B_POST_LAUNCH();
    }
    return;
  } else {
    dim3 grid((N + SBN - 1) / SBN, (unsigned)((M + SBM - 1) / SBM));
    simt_gemm_kernel<<<grid, 256, 0, st>>>(A, lda, M, W, ldw, N, K, ep);  // plain launch: PDL broke parity here
  }
  F5B_POST_LAUNCH();
}

// ---------------------------------------------------------------------------------------------
// Grouped Conv1d (ConvPositionEmbedding, reference modules.py:170-196), fp32.
//   x [batches, rows, channels]; w_r [groups][taps][cg_in][cg_out]; block = 32 rows x one group.
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
simt_gconv_kernel(const float* __restrict__ x, int rows, int channels, const float* __restrict__ w_r, int taps,
                  int pad, int cg, const EpiParams ep) {
  extern __shared__ float gc_smem[];
  const int span = 32 + taps - 1;
  float* Xs = gc_smem;              // [span][cg]
  float* Ws = gc_smem + span * cg;  // [cg][cg]
  const int g = blockIdx.y, b = blockIdx.z, t0 = blockIdx.x * 32;
  const float* xb = x + (long long)b * rows * channels + g * cg;
  for (int idx = threadIdx.x; idx < span * cg; idx += 256) {
    int r = idx / cg, c = idx - r * cg;
    int t = t0 + r - pad;
    Xs[idx] = (t >= 0 && t < rows) ? xb[(long long)t * channels + c] : 0.0f;
  }
  const int items = 32 * (cg / 4);  // (row, 4 output channels)
  float acc[2][4] = {{0, 0, 0, 0}, {0, 0, 0, 0}};
  for (int tap = 0; tap < taps; ++tap) {
    __syncthreads();
    const float* wt = w_r + ((long long)g * taps + tap) * cg * cg;
    for (int idx = threadIdx.x; idx < cg * cg; idx += 256) Ws[idx] = wt[idx];
    __syncthreads();
#pragma unroll
    for (int it = 0; it < 2; ++it) {
      int item = threadIdx.x + it * 256;
      if (item < items) {
        int r = item / (cg / 4), c4 = (item - r * (cg / 4)) * 4;
        const float* xr = Xs + (r + tap) * cg;
        for (int ci = 0; ci < cg; ++ci) {
          float xv = xr[ci];
          float4 w = *reinterpret_cast<const float4*>(Ws + ci * cg + c4);
          acc[it][0] = fmaf(xv, w.x, acc[it][0]);
          acc[it][1] = fmaf(xv, w.y, acc[it][1]);
          acc[it][2] = fmaf(xv, w.z, acc[it][2]);
          acc[it][3] = fmaf(xv, w.w, acc[it][3]);
        }
      }
    }
  }
#pragma unroll
  for (int it = 0; it < 2; ++it) {
    int item = threadIdx.x + it * 256;
    if (item < items) {
      int r = item / (cg / 4), c4 = (item - r * (cg / 4)) * 4;
      int t = t0 + r;
      if (t < rows) epi_apply<4>(ep, (long long)b * rows + t, g * cg + c4, acc[it]);
    }
  }
}

void simt_group_conv(cudaStream_t st, const float* x, int batches, int rows, int channels, const float* w_r, int taps,
                     int pad, int groups, const EpiParams& ep) {
  const int cg = channels / groups;
  F5B_REQUIRE(cg * groups == channels && cg % 4 == 0 && cg <= 64, "grouped conv: channels/group must be a multiple of 4, <= 64");
  const int smem = ((32 + taps - 1) * cg + cg * cg) * (int)sizeof(float);
  dim3 grid((rows + 31) / 32, groups, batches);
  simt_gconv_kernel<<<grid, 256, smem, st>>>(x, rows, channels, w_r, taps, pad, cg, ep);
  F5B_POST_LAUNCH();
}

// ---------------------------------------------------------------------------------------------
// fp32 attention (reference modules.py:459-544, torch backend): non-causal, head_dim 64, keys >= kv_len masked.
// One thread per query row, K/V tiles of 32 keys broadcast from smem, online softmax per tile.
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(128)
simt_attn_kernel(const float* __restrict__ qkv, float* __restrict__ out, int rows, int heads,
                 const int* __restrict__ kv_len, float scale) {
  __shared__ float Ks[32][64];
  __shared__ float Vs[32][64];
  const int b = blockIdx.z, h = blockIdx.y;
  const int inner = heads * 64;
  const long long ld = 3LL * inner;
  const int r = blockIdx.x * 128 + threadIdx.x;
  const bool live = r < rows;
  const int L = kv_len ? kv_len[b] : rows;
  const float* base = qkv + (long long)b * rows * ld;

  float q[64], o[64];
  {
    const float* qp = base + (long long)(live ? r : 0) * ld + h * 64;
#pragma unroll
    for (int d = 0; d < 64; d += 4) {
      float4 v = *reinterpret_cast<const float4*>(qp + d);
      q[d] = v.x * scale; q[d + 1] = v.y * scale; q[d + 2] = v.z * scale; q[d + 3] = v.w * scale;
    }
#pragma unroll
    for (int d = 0; d < 64; ++d) o[d] = 0.0f;
  }
  float mx = -INFINITY, l = 0.0f;

  for (int kv0 = 0; kv0 < L; kv0 += 32) {
    __syncthreads();
    for (int idx = threadIdx.x; idx < 32 * 16; idx += 128) {
      int j = idx >> 4, d4 = (idx & 15) * 4;
      int kv = kv0 + j;
      float4 kq = make_float4(0, 0, 0, 0), vq = kq;
      if (kv < L) {
        const float* rowp = base + (long long)kv * ld + h * 64 + d4;
        kq = *reinterpret_cast<const float4*>(rowp + inner);
        vq = *reinterpret_cast<const float4*>(rowp + 2 * inner);
      }
      *reinterpret_cast<float4*>(&Ks[j][d4]) = kq;
      *reinterpret_cast<float4*>(&Vs[j][d4]) = vq;
    }
    __syncthreads();
    float s[32];
    float cmax = -INFINITY;
#pragma unroll
    for (int j = 0; j < 32; ++j) {
      float a0 = 0.f, a1 = 0.f, a2 = 0.f, a3 = 0.f;
#pragma unroll
      for (int d = 0; d < 64; d += 4) {
        float4 kq = *reinterpret_cast<const float4*>(&Ks[j][d]);
        a0 = fmaf(q[d], kq.x, a0);
        a1 = fmaf(q[d + 1], kq.y, a1);
        a2 = fmaf(q[d + 2], kq.z, a2);
        a3 = fmaf(q[d + 3], kq.w, a3);
      }
      float sv = (a0 + a1) + (a2 + a3);
      s[j] = (kv0 + j < L) ? sv : -INFINITY;
      cmax = fmaxf(cmax, s[j]);
    }
    const float mnew = fmaxf(mx, cmax);
    const float alpha = expf(mx - mnew);  // mx = -inf on the first tile -> 0
    l *= alpha;
#pragma unroll
    for (int d = 0; d < 64; ++d) o[d] *= alpha;
#pragma unroll
    for (int j = 0; j < 32; ++j) {
      float p = expf(s[j] - mnew);
      l += p;
#pragma unroll
      for (int d = 0; d < 64; d += 4) {
        float4 vq = *reinterpret_cast<const float4*>(&Vs[j][d]);
        o[d] = fmaf(p, vq.x, o[d]);
        o[d + 1] = fmaf(p, vq.y, o[d + 1]);
        o[d + 2] = fmaf(p, vq.z, o[d + 2]);
        o[d + 3] = fmaf(p, vq.w, o[d + 3]);
      }
    }
    mx = mnew;
  }
  if (live) {
    const float inv = 1.0f / l;
    float* op = out + ((long long)b * rows + r) * inner + h * 64;
#pragma unroll
    for (int d = 0; d < 64; d += 4)
      *reinterpret_cast<float4*>(op + d) = make_float4(o[d] * inv, o[d + 1] * inv, o[d + 2] * inv, o[d + 3] * inv);
  }
}

void simt_init_kernels() {
  F5B_CHECK_CUDA(cudaFuncSetAttribute(skinny_gemm_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, SK_SMEM));
}

void simt_attention(cudaStream_t st, const float* qkv, float* out, int batches, int rows, int heads, const int* kv_len,
                    float scale) {
  dim3 grid((rows + 127) / 128, heads, batches);
  simt_attn_kernel<<<grid, 128, 0, st>>>(qkv, out, rows, heads, kv_len, scale);
  F5B_POST_LAUNCH();
}

}  // namespace f5b
