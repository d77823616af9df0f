// Backward kernels of the flow-matching training step (reference src/f5_tts/model/cfm.py:231-302 differentiated by
// trainer.py:384 `accelerator.backward(loss)`): everything needed to push the loss gradient through proj_out, the final
// AdaLN and a DiTBlock (modules.py:683-697) that is not a tensor-core contraction.  The fp32 CUDA-core GEMM / attention
// kernels here are the parity tier of the backward pass (checked against torch.autograd over the unmodified reference,
// tests/test_gpu_backward.py); in 16-bit mode the contractions run on tc_gemm / tc_attn / tc_attn_bwd and this file
// supplies the staging (cast, transpose, prep_dy16), the reductions and the elementwise derivatives around them.
// Deterministic: no atomics, fixed reduction order.
#include "train.h"

namespace f5b {

static inline unsigned cdiv(long long a, long long b) { return (unsigned)((a + b - 1) / b); }

// ---------------------------------------------------------------------------------------------
// C[i, j] (+)= sum_k A(i, k) B(k, j) with arbitrary element strides: every dgrad / wgrad form is one call
//   dX = dY W        : A = dY [M,N] (sa = N,1), B = W [N,K] (sb = K,1)
//   dW = dY^T X      : A = dY^T    (sa = 1,N), B = X [M,K] (sb = K,1)
// 64x64x16 tiles, 256 threads, 4x4 per thread.
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
sgemm_strided_kernel(const float* __restrict__ A, long long sa0, long long sa1, const float* __restrict__ B,
                     long long sb0, long long sb1, float* __restrict__ C, long long ldc, int M, int N, int K,
                     int accumulate, int k_slice) {
  __shared__ float As[16][64 + 1], Bs[16][64 + 1];
  const int tid = threadIdx.x, tx = tid & 15, ty = tid >> 4;
  const int m0 = blockIdx.y * 64, n0 = blockIdx.x * 64;
  float acc[4][4] = {};
  int kbeg = 0;
  if (k_slice > 0) {  // split-K: slice blockIdx.z of the contraction into its own [M, N] partial (C = scratch, ldc = N)
    kbeg = blockIdx.z * k_slice;
    K = min(K, kbeg + k_slice);
    C += (long long)blockIdx.z * M * N;
  }
  for (int k0 = kbeg; k0 < K; k0 += 16) {
    for (int e = tid; e < 64 * 16; e += 256) {
      // contiguous index first along whichever dimension is unit-stride in memory
      int i, k;
      if (sa1 == 1) { k = e & 15; i = e >> 4; } else { i = e & 63; k = e >> 6; }
      As[k][i] = (m0 + i < M && k0 + k < K) ? A[(long long)(m0 + i) * sa0 + (long long)(k0 + k) * sa1] : 0.0f;
      int j, kb;
      if (sb1 == 1) { j = e & 63; kb = e >> 6; } else { kb = e & 15; j = e >> 4; }
      Bs[kb][j] = (n0 + j < N && k0 + kb < K) ? B[(long long)(k0 + kb) * sb0 + (long long)(n0 + j) * sb1] : 0.0f;
    }
    __syncthreads();
#pragma unroll
    for (int k = 0; k < 16; ++k) {
      float a[4], b[4];
#pragma unroll
      for (int i = 0; i < 4; ++i) a[i] = As[k][ty * 4 + i];
#pragma unroll
      for (int j = 0; j < 4; ++j) b[j] = Bs[k][tx * 4 + j];
#pragma unroll
      for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(a[i], b[j], acc[i][j]);
    }
    __syncthreads();
  }
#pragma unroll
  for (int i = 0; i < 4; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int m = m0 + ty * 4 + i, n = n0 + tx * 4 + j;
      if (m < M && n < N) {
        float* c = C + (long long)m * ldc + n;
        *c = accumulate ? *c + acc[i][j] : acc[i][j];
      }
    }
}

// Skinny product with a long contraction (ds += dmod [B, 6d] W_ada [6d, d]: 40 x 1024 x 6144): 16 CTAs walking 384
// k-blocks each took 0.55 ms.  K is cut into `splits` slices (blockIdx.z), every slice writes its partial into the
// scratch, a second kernel adds the slices in order (deterministic) into C.
__global__ void __launch_bounds__(256)
splitk_reduce_kernel(const float* __restrict__ part, int splits, long long mn, int N, float* __restrict__ C, long long ldc,
                     int accumulate) {
  const long long i = (long long)blockIdx.x * 256 + threadIdx.x;
  if (i >= mn) return;
  float acc = 0.0f;
  for (int s = 0; s < splits; ++s) acc += part[(long long)s * mn + i];
  float* c = C + (i / N) * ldc + (i % N);
  *c = accumulate ? *c + acc : acc;
}
static thread_local float* g_splitk_scratch = nullptr;
static thread_local size_t g_splitk_scratch_floats = 0;
void sgemm_set_scratch(float* p, size_t floats) {
  g_splitk_scratch = p;
  g_splitk_scratch_floats = floats;
}
void sgemm_strided(cudaStream_t st, const float* A, long long sa0, long long sa1, const float* B, long long sb0,
                   long long sb1, float* C, long long ldc, int M, int N, int K, bool accumulate) {
  if (M <= 64 && K >= 1024 && g_splitk_scratch != nullptr) {
    int splits = K / 96;  // slices of >= 96 (a multiple of the 16-deep k-block)
    if (splits > 64) splits = 64;
    while (splits > 1 && (size_t)splits * M * N > g_splitk_scratch_floats) --splits;
    if (splits > 1) {
      const int k_slice = ((K + splits - 1) / splits + 15) / 16 * 16;
      splits = (K + k_slice - 1) / k_slice;
      sgemm_strided_kernel<<<dim3(cdiv(N, 64), cdiv(M, 64), splits), 256, 0, st>>>(A, sa0, sa1, B, sb0, sb1,
                                                                                  g_splitk_scratch, N, M, N, K, 0, k_slice);
      F5B_POST_LAUNCH();
      splitk_reduce_kernel<<<cdiv((long long)M * N, 256), 256, 0, st>>>(g_splitk_scratch, splits, (long long)M * N, N, C,
                                                                       ldc, accumulate ? 1 : 0);
      F5B_POST_LAUNCH();
      return;
    }
  }
  sgemm_strided_kernel<<<dim3(cdiv(N, 64), cdiv(M, 64)), 256, 0, st>>>(A, sa0, sa1, B, sb0, sb1, C, ldc, M, N, K,
                                                                          accumulate ? 1 : 0, 0);
  F5B_POST_LAUNCH();
}

// out[b, j] = sum over the rows r of batch element b of X[r, j] * (Y ? Y[r, j] : 1); one thread per (b, j), rows in order
__global__ void __launch_bounds__(256)
colsum_batch_kernel(const float* __restrict__ X, const float* __restrict__ Y, int rows_per_batch, int cols,
                    float* __restrict__ out, long long out_bstride) {
  const int j = blockIdx.x * 256 + threadIdx.x, b = blockIdx.y;
  if (j >= cols) return;
  const float* x = X + (long long)b * rows_per_batch * cols + j;
  const float* y = Y ? Y + (long long)b * rows_per_batch * cols + j : nullptr;
  float s = 0.0f;
  for (int r = 0; r < rows_per_batch; ++r) s += y ? x[(long long)r * cols] * y[(long long)r * cols] : x[(long long)r * cols];
  out[(long long)b * out_bstride + j] = s;
}
// two-level form for one long batch element (the bias / affine gradients sum over ALL rows): chunk sums, then the chunks
// in order -- still one owner per output, no atomics
__global__ void __launch_bounds__(256)
colsum_chunk_kernel(const float* __restrict__ X, const float* __restrict__ Y, int rows, int cols, int chunk,
                    float* __restrict__ part) {
  const int j = blockIdx.x * 256 + threadIdx.x, s = blockIdx.y;
  if (j >= cols) return;
  const int r0 = s * chunk, r1 = min(rows, r0 + chunk);
  float acc = 0.0f;
  for (int r = r0; r < r1; ++r) {
    const long long i = (long long)r * cols + j;
    acc += Y ? X[i] * Y[i] : X[i];
  }
  part[(long long)s * cols + j] = acc;
}
__global__ void __launch_bounds__(256)
colsum_final_kernel(const float* __restrict__ part, int chunks, int cols, float* __restrict__ out) {
  const int j = blockIdx.x * 256 + threadIdx.x;
  if (j >= cols) return;
  float acc = 0.0f;
  for (int s = 0; s < chunks; ++s) acc += part[(long long)s * cols + j];
  out[j] = acc;
}
// 16-byte form of both kernels above: a CTA covers 256 columns (64 float4 lanes) of one segment of rows (a batch element
// or a chunk) with FOUR row groups (rows r0 + g, r0 + g + 4, ...), each thread keeping two independent sums in flight;
// the groups are added in order through shared memory -- one owner per output, fixed order, 4x the loads in flight.
__global__ void __launch_bounds__(256)
colsum_seg4_kernel(const float* __restrict__ X, const float* __restrict__ Y, int seg_rows, int rows_total, int cols,
                   float* __restrict__ out, long long out_sstride) {
  __shared__ float4 red[4][64];
  const int q = threadIdx.x & 63, g = threadIdx.x >> 6, s = blockIdx.y;
  const int c = (blockIdx.x * 64 + q) * 4;
  const long long r0 = (long long)s * seg_rows;
  const long long r1 = min((long long)rows_total, r0 + seg_rows);
  float4 a0 = make_float4(0.f, 0.f, 0.f, 0.f), a1 = a0;
  if (c < cols) {
    long long r = r0 + g;
    for (; r + 4 < r1; r += 8) {
      const float4 x0 = *reinterpret_cast<const float4*>(X + r * cols + c);
      const float4 x1 = *reinterpret_cast<const float4*>(X + (r + 4) * cols + c);
      if (Y) {
        const float4 y0 = *reinterpret_cast<const float4*>(Y + r * cols + c);
        const float4 y1 = *reinterpret_cast<const float4*>(Y + (r + 4) * cols + c);
        a0.x = fmaf(x0.x, y0.x, a0.x); a0.y = fmaf(x0.y, y0.y, a0.y); a0.z = fmaf(x0.z, y0.z, a0.z); a0.w = fmaf(x0.w, y0.w, a0.w);
        a1.x = fmaf(x1.x, y1.x, a1.x); a1.y = fmaf(x1.y, y1.y, a1.y); a1.z = fmaf(x1.z, y1.z, a1.z); a1.w = fmaf(x1.w, y1.w, a1.w);
      } else {
        a0.x += x0.x; a0.y += x0.y; a0.z += x0.z; a0.w += x0.w;
        a1.x += x1.x; a1.y += x1.y; a1.z += x1.z; a1.w += x1.w;
      }
    }
    if (r < r1) {
      const float4 x0 = *reinterpret_cast<const float4*>(X + r * cols + c);
      if (Y) {
        const float4 y0 = *reinterpret_cast<const float4*>(Y + r * cols + c);
        a0.x = fmaf(x0.x, y0.x, a0.x); a0.y = fmaf(x0.y, y0.y, a0.y); a0.z = fmaf(x0.z, y0.z, a0.z); a0.w = fmaf(x0.w, y0.w, a0.w);
      } else {
        a0.x += x0.x; a0.y += x0.y; a0.z += x0.z; a0.w += x0.w;
      }
    }
  }
  red[g][q] = make_float4(a0.x + a1.x, a0.y + a1.y, a0.z + a1.z, a0.w + a1.w);
  __syncthreads();
  if (g == 0 && c < cols) {
    float4 t = red[0][q];
#pragma unroll
    for (int k = 1; k < 4; ++k) {
      t.x += red[k][q].x; t.y += red[k][q].y; t.z += red[k][q].z; t.w += red[k][q].w;
    }
    float* o = out + (long long)s * out_sstride + c;
    o[0] = t.x; o[1] = t.y; o[2] = t.z; o[3] = t.w;
  }
}
static thread_local float* g_colsum_scratch = nullptr;
static thread_local size_t g_colsum_scratch_floats = 0;
void colsum_set_scratch(float* p, size_t floats) {
  g_colsum_scratch = p;
  g_colsum_scratch_floats = floats;
}
float* colsum_scratch(size_t* floats) {
  *floats = g_colsum_scratch_floats;
  return g_colsum_scratch;
}
// Both AdaLN sums of one upstream gradient in one pass: out_xy[b, j] = sum_r X Y, out_x[b, j] = sum_r X (per batch
// element) -- d scale and d shift of a modulated LayerNorm read the same X; the pair of colsum_batch calls read it twice.
__global__ void __launch_bounds__(256)
colsum_pair4_kernel(const float* __restrict__ X, const float* __restrict__ Y, int seg_rows, int cols,
                    float* __restrict__ out_xy, float* __restrict__ out_x, long long out_sstride) {
  __shared__ float4 red[2][4][64];
  const int q = threadIdx.x & 63, g = threadIdx.x >> 6, s = blockIdx.y;
  const int c = (blockIdx.x * 64 + q) * 4;
  const long long r0 = (long long)s * seg_rows, r1 = r0 + seg_rows;
  float4 a = make_float4(0.f, 0.f, 0.f, 0.f), b = a;
  if (c < cols) {
    for (long long r = r0 + g; r < r1; r += 4) {
      const float4 x = *reinterpret_cast<const float4*>(X + r * cols + c);
      const float4 y = *reinterpret_cast<const float4*>(Y + r * cols + c);
      a.x = fmaf(x.x, y.x, a.x); a.y = fmaf(x.y, y.y, a.y); a.z = fmaf(x.z, y.z, a.z); a.w = fmaf(x.w, y.w, a.w);
      b.x += x.x; b.y += x.y; b.z += x.z; b.w += x.w;
    }
  }
  red[0][g][q] = a;
  red[1][g][q] = b;
  __syncthreads();
  if (g < 2 && c < cols) {  // group 0 finishes the X Y sums, group 1 the X sums: the four row groups in order
    float4 t = red[g][0][q];
#pragma unroll
    for (int k = 1; k < 4; ++k) {
      t.x += red[g][k][q].x; t.y += red[g][k][q].y; t.z += red[g][k][q].z; t.w += red[g][k][q].w;
    }
    float* o = (g == 0 ? out_xy : out_x) + (long long)s * out_sstride + c;
    o[0] = t.x; o[1] = t.y; o[2] = t.z; o[3] = t.w;
  }
}
void colsum_batch_pair(cudaStream_t st, const float* X, const float* Y, int batches, int rows_per_batch, int cols,
                       float* out_xy, float* out_x, long long out_bstride) {
  if (cols % 4 == 0 && (reinterpret_cast<uintptr_t>(X) & 15) == 0 && (reinterpret_cast<uintptr_t>(Y) & 15) == 0) {
    colsum_pair4_kernel<<<dim3(cdiv(cols, 256), batches), 256, 0, st>>>(X, Y, rows_per_batch, cols, out_xy, out_x,
                                                                       out_bstride);
    F5B_POST_LAUNCH();
    return;
  }
  colsum_batch(st, X, Y, batches, rows_per_batch, cols, out_xy, out_bstride);
  colsum_batch(st, X, nullptr, batches, rows_per_batch, cols, out_x, out_bstride);
}
void colsum_batch(cudaStream_t st, const float* X, const float* Y, int batches, int rows_per_batch, int cols, float* out,
                  long long out_bstride) {
  constexpr int kChunks = 128;
  const bool vec = cols % 4 == 0 && (reinterpret_cast<uintptr_t>(X) & 15) == 0 &&
                   (Y == nullptr || (reinterpret_cast<uintptr_t>(Y) & 15) == 0);
  if (batches == 1 && rows_per_batch >= 4096 && g_colsum_scratch != nullptr &&
      (size_t)kChunks * cols <= g_colsum_scratch_floats) {
    const int chunk = (rows_per_batch + kChunks - 1) / kChunks;
    if (vec) {
      const int chunks = (rows_per_batch + chunk - 1) / chunk;
      colsum_seg4_kernel<<<dim3(cdiv(cols, 256), chunks), 256, 0, st>>>(X, Y, chunk, rows_per_batch, cols,
                                                                       g_colsum_scratch, cols);
      F5B_POST_LAUNCH();
      colsum_final_kernel<<<cdiv(cols, 256), 256, 0, st>>>(g_colsum_scratch, chunks, cols, out);
      F5B_POST_LAUNCH();
      return;
    }
    colsum_chunk_kernel<<<dim3(cdiv(cols, 256), kChunks), 256, 0, st>>>(X, Y, rows_per_batch, cols, chunk, g_colsum_scratch);
    F5B_POST_LAUNCH();
    colsum_final_kernel<<<cdiv(cols, 256), 256, 0, st>>>(g_colsum_scratch, kChunks, cols, out);
    F5B_POST_LAUNCH();
    return;
  }
  if (vec && rows_per_batch >= 64) {
    colsum_seg4_kernel<<<dim3(cdiv(cols, 256), batches), 256, 0, st>>>(X, Y, rows_per_batch, batches * rows_per_batch,
                                                                      cols, out, out_bstride);
    F5B_POST_LAUNCH();
    return;
  }
  colsum_batch_kernel<<<dim3(cdiv(cols, 256), batches), 256, 0, st>>>(X, Y, rows_per_batch, cols, out, out_bstride);
  F5B_POST_LAUNCH();
}

// ---------------------------------------------------------------------------------------------
// LayerNorm (no affine, eps) + modulation, forward with saved statistics, and its backward
//   y = xhat * (1 + scale[b]) + shift[b];   g = dy * (1 + scale[b]);   dx (+)= rstd * (g - mean(g) - xhat * mean(g xhat))
// one warp per row
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
ln_mod_fwd_save_kernel(const float* __restrict__ x, long long M, int d, const float* __restrict__ scale,
                       const float* __restrict__ shift, long long vec_bstride, int rows_per_batch, float eps,
                       float* __restrict__ y, float* __restrict__ xhat, float* __restrict__ rstd_out) {
  const long long m = (long long)blockIdx.x * 8 + (threadIdx.x >> 5);
  const int lane = threadIdx.x & 31;
  if (m >= M) return;
  const float* xr = x + m * d;
  float s = 0.0f;
  for (int c = lane; c < d; c += 32) s += xr[c];
  const float mean = warp_sum(s) / d;
  float q = 0.0f;
  for (int c = lane; c < d; c += 32) {
    const float t = xr[c] - mean;
    q += t * t;
  }
  const float rstd = rsqrtf(warp_sum(q) / d + eps);
  const long long b = m / rows_per_batch;
  const float* sc = scale + b * vec_bstride;
  const float* sh = shift + b * vec_bstride;
  for (int c = lane; c < d; c += 32) {
    const float h = (xr[c] - mean) * rstd;
    xhat[m * d + c] = h;
    y[m * d + c] = h * (1.0f + sc[c]) + sh[c];
  }
  if (lane == 0) rstd_out[m] = rstd;
}
void ln_mod_fwd_save(cudaStream_t st, const float* x, long long M, int d, const float* scale, const float* shift,
                     long long vec_bstride, int rows_per_batch, float eps, float* y, float* xhat, float* rstd) {
  ln_mod_fwd_save_kernel<<<cdiv(M, 8), 256, 0, st>>>(x, M, d, scale, shift, vec_bstride, rows_per_batch, eps, y, xhat, rstd);
  F5B_POST_LAUNCH();
}

static inline bool al16f(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15) == 0; }
__global__ void __launch_bounds__(256)
ln_mod_bwd_kernel(const float* __restrict__ dy, const float* __restrict__ xhat, const float* __restrict__ rstd,
                  long long M, int d, const float* __restrict__ scale, long long vec_bstride, int rows_per_batch,
                  float* __restrict__ dx, int accumulate) {
  const long long m = (long long)blockIdx.x * 8 + (threadIdx.x >> 5);
  const int lane = threadIdx.x & 31;
  if (m >= M) return;
  const float* sc = scale + (m / rows_per_batch) * vec_bstride;
  const float* g = dy + m * d;
  const float* h = xhat + m * d;
  float s1 = 0.0f, s2 = 0.0f;
  for (int c = lane; c < d; c += 32) {
    const float gv = g[c] * (1.0f + sc[c]);
    s1 += gv;
    s2 += gv * h[c];
  }
  const float m1 = warp_sum(s1) / d, m2 = warp_sum(s2) / d, r = rstd[m];
  for (int c = lane; c < d; c += 32) {
    const float v = r * (g[c] * (1.0f + sc[c]) - m1 - h[c] * m2);
    dx[m * d + c] = accumulate ? dx[m * d + c] + v : v;
  }
}
// Register-resident form for d = 128 k <= 1024 (the row of dy and of xhat is read ONCE, 16 bytes per lane and access):
//   dx = (resid ? resid : 0) + rstd * (g - mean(g) - xhat * mean(g xhat)),  g = dy * (1 + scale[b])
// `resid` may alias dx (accumulate in place) or be another buffer (the residual branch's gradient: no copy needed first).
__global__ void __launch_bounds__(256)
ln_mod_bwd4_kernel(const float* __restrict__ dy, const float* __restrict__ xhat, const float* __restrict__ rstd,
                   long long M, int d, const float* __restrict__ scale, long long vec_bstride, int rows_per_batch,
                   const float* resid, float* dx) {
  const long long m = (long long)blockIdx.x * 8 + (threadIdx.x >> 5);
  const int lane = threadIdx.x & 31;
  if (m >= M) return;
  const float* sc = scale + (m / rows_per_batch) * vec_bstride;
  float4 g[8], h[8];
  float s1 = 0.0f, s2 = 0.0f;
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    const int c = i * 128 + lane * 4;
    if (c < d) {
      const float4 a = *reinterpret_cast<const float4*>(dy + m * d + c);
      const float4 s = *reinterpret_cast<const float4*>(sc + c);
      h[i] = *reinterpret_cast<const float4*>(xhat + m * d + c);
      g[i] = make_float4(a.x * (1.0f + s.x), a.y * (1.0f + s.y), a.z * (1.0f + s.z), a.w * (1.0f + s.w));
      s1 += (g[i].x + g[i].y) + (g[i].z + g[i].w);
      s2 += (g[i].x * h[i].x + g[i].y * h[i].y) + (g[i].z * h[i].z + g[i].w * h[i].w);
    }
  }
  const float m1 = warp_sum(s1) / d, m2 = warp_sum(s2) / d, r = rstd[m];
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    const int c = i * 128 + lane * 4;
    if (c < d) {
      float4 v = make_float4(r * (g[i].x - m1 - h[i].x * m2), r * (g[i].y - m1 - h[i].y * m2),
                             r * (g[i].z - m1 - h[i].z * m2), r * (g[i].w - m1 - h[i].w * m2));
      if (resid != nullptr) {
        const float4 b = *reinterpret_cast<const float4*>(resid + m * d + c);
        v.x += b.x; v.y += b.y; v.z += b.z; v.w += b.w;
      }
      *reinterpret_cast<float4*>(dx + m * d + c) = v;
    }
  }
}
void ln_mod_bwd_resid(cudaStream_t st, const float* dy, const float* xhat, const float* rstd, long long M, int d,
                      const float* scale, long long vec_bstride, int rows_per_batch, const float* resid, float* dx) {
  if (d % 128 == 0 && d <= 1024 && vec_bstride % 4 == 0 && al16f(dy) && al16f(xhat) && al16f(scale) && al16f(dx) &&
      (resid == nullptr || al16f(resid))) {
    ln_mod_bwd4_kernel<<<cdiv(M, 8), 256, 0, st>>>(dy, xhat, rstd, M, d, scale, vec_bstride, rows_per_batch, resid, dx);
    F5B_POST_LAUNCH();
    return;
  }
  if (resid != nullptr && resid != dx)
    F5B_CHECK_CUDA(cudaMemcpyAsync(dx, resid, (size_t)M * d * 4, cudaMemcpyDeviceToDevice, st));
  ln_mod_bwd_kernel<<<cdiv(M, 8), 256, 0, st>>>(dy, xhat, rstd, M, d, scale, vec_bstride, rows_per_batch, dx, resid ? 1 : 0);
  F5B_POST_LAUNCH();
}
void ln_mod_bwd(cudaStream_t st, const float* dy, const float* xhat, const float* rstd, long long M, int d,
                const float* scale, long long vec_bstride, int rows_per_batch, float* dx, bool accumulate) {
  ln_mod_bwd_resid(st, dy, xhat, rstd, M, d, scale, vec_bstride, rows_per_batch, accumulate ? dx : nullptr, dx);
}

// ---------------------------------------------------------------------------------------------
// elementwise pieces
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
gelu_tanh_fwd_kernel(const float* __restrict__ u, float* __restrict__ h, long long n) {
  const long long i = (long long)blockIdx.x * 256 + threadIdx.x;
  if (i < n) h[i] = act_apply(ACT_GELU_TANH, u[i]);
}
// du = dh * d/du [0.5 u (1 + tanh(k (u + c u^3)))]
__global__ void __launch_bounds__(256)
gelu_tanh_bwd_kernel(const float* __restrict__ u, const float* __restrict__ dh, float* __restrict__ du, long long n) {
  const long long i = (long long)blockIdx.x * 256 + threadIdx.x;
  if (i >= n) return;
  const float x = u[i], k = 0.7978845608028654f, c = 0.044715f;
  const float t = tanhf(k * (x + c * x * x * x));
  du[i] = dh[i] * (0.5f * (1.0f + t) + 0.5f * x * (1.0f - t * t) * k * (1.0f + 3.0f * c * x * x));
}
__device__ __forceinline__ float gelu_tanh_grad(float x, float dh) {
  const float k = 0.7978845608028654f, c = 0.044715f;
  const float t = tanhf(k * (x + c * x * x * x));
  return dh * (0.5f * (1.0f + t) + 0.5f * x * (1.0f - t * t) * k * (1.0f + 3.0f * c * x * x));
}
// four elements per thread (16-byte accesses): the scalar forms above ran at a third of the HBM rate
__global__ void __launch_bounds__(256)
gelu_tanh_fwd4_kernel(const float4* __restrict__ u, float4* __restrict__ h, long long n4) {
  const long long i = (long long)blockIdx.x * 256 + threadIdx.x;
  if (i >= n4) return;
  const float4 x = u[i];
  h[i] = make_float4(act_apply(ACT_GELU_TANH, x.x), act_apply(ACT_GELU_TANH, x.y), act_apply(ACT_GELU_TANH, x.z),
                     act_apply(ACT_GELU_TANH, x.w));
}
__global__ void __launch_bounds__(256)
gelu_tanh_bwd4_kernel(const float4* __restrict__ u, const float4* __restrict__ dh, float4* __restrict__ du, long long n4) {
  const long long i = (long long)blockIdx.x * 256 + threadIdx.x;
  if (i >= n4) return;
  const float4 x = u[i], g = dh[i];
  du[i] = make_float4(gelu_tanh_grad(x.x, g.x), gelu_tanh_grad(x.y, g.y), gelu_tanh_grad(x.z, g.z), gelu_tanh_grad(x.w, g.w));
}
static inline bool al16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15) == 0; }
void gelu_tanh_fwd(cudaStream_t st, const float* u, float* h, long long n) {
  if (n % 4 == 0 && al16(u) && al16(h))
    gelu_tanh_fwd4_kernel<<<cdiv(n / 4, 256), 256, 0, st>>>(reinterpret_cast<const float4*>(u), reinterpret_cast<float4*>(h), n / 4);
  else
    gelu_tanh_fwd_kernel<<<cdiv(n, 256), 256, 0, st>>>(u, h, n);
  F5B_POST_LAUNCH();
}
void gelu_tanh_bwd(cudaStream_t st, const float* u, const float* dh, float* du, long long n) {
  if (n % 4 == 0 && al16(u) && al16(dh) && al16(du))
    gelu_tanh_bwd4_kernel<<<cdiv(n / 4, 256), 256, 0, st>>>(reinterpret_cast<const float4*>(u),
                                                          reinterpret_cast<const float4*>(dh),
                                                          reinterpret_cast<float4*>(du), n / 4);
  else
    gelu_tanh_bwd_kernel<<<cdiv(n, 256), 256, 0, st>>>(u, dh, du, n);
  F5B_POST_LAUNCH();
}

// dz = dy * d/dz [z sigmoid(z)]
__global__ void __launch_bounds__(256)
silu_bwd_kernel(const float* __restrict__ z, const float* __restrict__ dy, float* __restrict__ dz, long long n) {
  const long long i = (long long)blockIdx.x * 256 + threadIdx.x;
  if (i >= n) return;
  const float s = 1.0f / (1.0f + expf(-z[i]));
  dz[i] = dy[i] * s * (1.0f + z[i] * (1.0f - s));
}
void silu_bwd(cudaStream_t st, const float* z, const float* dy, float* dz, long long n) {
  silu_bwd_kernel<<<cdiv(n, 256), 256, 0, st>>>(z, dy, dz, n);
  F5B_POST_LAUNCH();
}

// out[m, :] = v[b(m), :] * x[m, :] (+ bias[:]) ; rows at or beyond row_len[b] give 0 when `zero_dead`
__global__ void __launch_bounds__(256)
rowscale_kernel(const float* __restrict__ x, const float* __restrict__ v, long long vec_bstride, int rows_per_batch,
                const int* __restrict__ row_len, int cols, long long total, float* __restrict__ out) {
  const long long i = (long long)blockIdx.x * 256 + threadIdx.x;
  if (i >= total) return;
  const long long m = i / cols;
  const int c = (int)(i - m * cols);
  const long long b = m / rows_per_batch;
  const int pos = (int)(m - b * rows_per_batch);
  const bool dead = row_len != nullptr && pos >= row_len[b];
  out[i] = dead ? 0.0f : (v ? v[b * vec_bstride + c] : 1.0f) * x[i];
}
__global__ void __launch_bounds__(256)
rowscale4_kernel(const float4* __restrict__ x, const float* __restrict__ v, long long vec_bstride, int rows_per_batch,
                 const int* __restrict__ row_len, int cols4, long long total4, float4* __restrict__ out) {
  const long long i = (long long)blockIdx.x * 256 + threadIdx.x;
  if (i >= total4) return;
  const long long m = i / cols4;
  const int c = (int)(i - m * cols4) * 4;
  const long long b = m / rows_per_batch;
  const int pos = (int)(m - b * rows_per_batch);
  if (row_len != nullptr && pos >= row_len[b]) {
    out[i] = make_float4(0.f, 0.f, 0.f, 0.f);
    return;
  }
  float4 a = x[i];
  if (v) {
    const float4 s = *reinterpret_cast<const float4*>(v + b * vec_bstride + c);
    a.x *= s.x; a.y *= s.y; a.z *= s.z; a.w *= s.w;
  }
  out[i] = a;
}
// out = base + v[b] * x   (the residual update x1 = x0 + gate * o in one pass instead of scale, copy, add)
__global__ void __launch_bounds__(256)
rowscale_add4_kernel(const float4* __restrict__ base, const float4* __restrict__ x, const float* __restrict__ v,
                     long long vec_bstride, int rows_per_batch, int cols4, long long total4, float4* __restrict__ out) {
  const long long i = (long long)blockIdx.x * 256 + threadIdx.x;
  if (i >= total4) return;
  const long long m = i / cols4;
  const int c = (int)(i - m * cols4) * 4;
  const long long b = m / rows_per_batch;
  const float4 s = *reinterpret_cast<const float4*>(v + b * vec_bstride + c);
  const float4 a = x[i], r = base[i];
  out[i] = make_float4(fmaf(s.x, a.x, r.x), fmaf(s.y, a.y, r.y), fmaf(s.z, a.z, r.z), fmaf(s.w, a.w, r.w));
}
__global__ void __launch_bounds__(256)
rowscale_add_kernel(const float* __restrict__ base, const float* __restrict__ x, const float* __restrict__ v,
                    long long vec_bstride, int rows_per_batch, int cols, long long total, float* __restrict__ out) {
  const long long i = (long long)blockIdx.x * 256 + threadIdx.x;
  if (i >= total) return;
  const long long m = i / cols;
  out[i] = fmaf(v[(m / rows_per_batch) * vec_bstride + (i - m * cols)], x[i], base[i]);
}
void rowscale_add(cudaStream_t st, const float* base, const float* x, const float* v, long long vec_bstride,
                  int rows_per_batch, long long M, int cols, float* out) {
  if (cols % 4 == 0 && vec_bstride % 4 == 0 && al16(base) && al16(x) && al16(out) && al16(v))
    rowscale_add4_kernel<<<cdiv(M * (cols / 4), 256), 256, 0, st>>>(reinterpret_cast<const float4*>(base),
                                                                   reinterpret_cast<const float4*>(x), v, vec_bstride,
                                                                   rows_per_batch, cols / 4, M * (cols / 4),
                                                                   reinterpret_cast<float4*>(out));
  else
    rowscale_add_kernel<<<cdiv(M * cols, 256), 256, 0, st>>>(base, x, v, vec_bstride, rows_per_batch, cols, M * cols, out);
  F5B_POST_LAUNCH();
}
void rowscale(cudaStream_t st, const float* x, const float* v, long long vec_bstride, int rows_per_batch,
              const int* row_len, long long M, int cols, float* out) {
  if (cols % 4 == 0 && vec_bstride % 4 == 0 && al16(x) && al16(out) && (v == nullptr || al16(v))) {
    rowscale4_kernel<<<cdiv(M * (cols / 4), 256), 256, 0, st>>>(reinterpret_cast<const float4*>(x), v, vec_bstride,
                                                               rows_per_batch, row_len, cols / 4, M * (cols / 4),
                                                               reinterpret_cast<float4*>(out));
    F5B_POST_LAUNCH();
    return;
  }
  rowscale_kernel<<<cdiv(M * cols, 256), 256, 0, st>>>(x, v, vec_bstride, rows_per_batch, row_len, cols, M * cols, out);
  F5B_POST_LAUNCH();
}

__global__ void __launch_bounds__(256)
axpy_kernel(float* __restrict__ y, const float* __restrict__ x, float a, long long n) {
  const long long i = (long long)blockIdx.x * 256 + threadIdx.x;
  if (i < n) y[i] += a * x[i];
}
__global__ void __launch_bounds__(256)
axpy4_kernel(float4* __restrict__ y, const float4* __restrict__ x, float a, long long n4) {
  const long long i = (long long)blockIdx.x * 256 + threadIdx.x;
  if (i >= n4) return;
  float4 r = y[i];
  const float4 v = x[i];
  r.x += a * v.x; r.y += a * v.y; r.z += a * v.z; r.w += a * v.w;
  y[i] = r;
}
void axpy(cudaStream_t st, float* y, const float* x, float a, long long n) {
  if (n % 4 == 0 && al16(y) && al16(x))
    axpy4_kernel<<<cdiv(n / 4, 256), 256, 0, st>>>(reinterpret_cast<float4*>(y), reinterpret_cast<const float4*>(x), a, n / 4);
  else
    axpy_kernel<<<cdiv(n, 256), 256, 0, st>>>(y, x, a, n);
  F5B_POST_LAUNCH();
}

__global__ void __launch_bounds__(256)
add_bias_kernel(float* __restrict__ y, const float* __restrict__ bias, int cols, long long total) {
  const long long i = (long long)blockIdx.x * 256 + threadIdx.x;
  if (i < total) y[i] += bias[i % cols];
}
void add_bias(cudaStream_t st, float* y, const float* bias, long long M, int cols) {
  add_bias_kernel<<<cdiv(M * cols, 256), 256, 0, st>>>(y, bias, cols, M * cols);
  F5B_POST_LAUNCH();
}

// RoPE on the q | k thirds of qkv [M, 3 * inner], interleaved pairs, heads < rope_heads; dir = +1 forward (R x),
// -1 transpose (R^T x: the backward of the rotation).  cs [rows_per_batch, 64] = (cos_j, sin_j) pairs.
__global__ void __launch_bounds__(256)
rope_qk_kernel(float* __restrict__ qkv, long long M, int inner, int rows_per_batch, int rope_heads,
               const float* __restrict__ cs, float dir) {
  const long long i = (long long)blockIdx.x * 256 + threadIdx.x;  // one thread per (row, q-or-k, pair)
  const int pairs = inner / 2;
  if (i >= M * 2 * pairs) return;
  const long long m = i / (2 * pairs);
  const int r = (int)(i - m * 2 * pairs), part = r / pairs, p = r - part * pairs;
  const int col = 2 * p, head = col >> 6;
  if (head >= rope_heads) return;
  const int pos = (int)(m % rows_per_batch);
  const float2 t = *reinterpret_cast<const float2*>(cs + (long long)pos * 64 + (col & 63));
  float* x = qkv + m * 3 * inner + (long long)part * inner + col;
  const float a = x[0], c = x[1], sn = dir * t.y;
  x[0] = a * t.x - c * sn;
  x[1] = c * t.x + a * sn;
}
// two pairs (one float4 of x, one float4 of the table) per thread
__global__ void __launch_bounds__(256)
rope_qk4_kernel(float* __restrict__ qkv, long long M, int inner, int rows_per_batch, int rope_heads,
                const float* __restrict__ cs, float dir) {
  const long long i = (long long)blockIdx.x * 256 + threadIdx.x;  // one thread per (row, q-or-k, 4 columns)
  const int quads = inner / 4;
  if (i >= M * 2 * quads) return;
  const long long m = i / (2 * quads);
  const int r = (int)(i - m * 2 * quads), part = r / quads, col = (r - part * quads) * 4;
  if ((col >> 6) >= rope_heads) return;
  const int pos = (int)(m % rows_per_batch);
  const float4 t = *reinterpret_cast<const float4*>(cs + (long long)pos * 64 + (col & 63));
  float4* x = reinterpret_cast<float4*>(qkv + m * 3 * inner + (long long)part * inner + col);
  const float4 v = *x;
  const float s0 = dir * t.y, s1 = dir * t.w;
  *x = make_float4(v.x * t.x - v.y * s0, v.y * t.x + v.x * s0, v.z * t.z - v.w * s1, v.w * t.z + v.z * s1);
}
void rope_qk(cudaStream_t st, float* qkv, long long M, int inner, int rows_per_batch, int rope_heads, const float* cs,
             float dir) {
  if (inner % 4 == 0 && al16(qkv) && al16(cs)) {
    rope_qk4_kernel<<<cdiv(M * (inner / 2), 256), 256, 0, st>>>(qkv, M, inner, rows_per_batch, rope_heads, cs, dir);
    F5B_POST_LAUNCH();
    return;
  }
  rope_qk_kernel<<<cdiv(M * inner, 256), 256, 0, st>>>(qkv, M, inner, rows_per_batch, rope_heads, cs, dir);
  F5B_POST_LAUNCH();
}

// masked-MSE gradient (cfm.py:298-300): loss = mean over the span frames x mel of (pred - (x1 - x0))^2
//   dpred = 2 (pred - flow) / count on span rows, 0 elsewhere; count = (#span rows) * mel, counted deterministically
__global__ void count_span_kernel(const unsigned char* __restrict__ span, long long rows, float* __restrict__ cnt) {
  __shared__ float part[256];
  float c = 0.0f;
  for (long long m = threadIdx.x; m < rows; m += 256) c += span[m] ? 1.0f : 0.0f;
  part[threadIdx.x] = c;
  __syncthreads();
  if (threadIdx.x == 0) {
    float t = 0.0f;
    for (int i = 0; i < 256; ++i) t += part[i];
    *cnt = t;
  }
}
__global__ void __launch_bounds__(256)
mse_grad_kernel(const float* __restrict__ pred, const float* __restrict__ x1, const float* __restrict__ x0,
                const unsigned char* __restrict__ span, const float* __restrict__ cnt, int mel, long long total,
                float* __restrict__ dpred) {
  const long long i = (long long)blockIdx.x * 256 + threadIdx.x;
  if (i >= total) return;
  const long long m = i / mel;
  dpred[i] = span[m] ? 2.0f * (pred[i] - (x1[i] - x0[i])) / (*cnt * mel) : 0.0f;
}
void mse_grad(cudaStream_t st, const float* pred, const float* x1, const float* x0, const unsigned char* span,
              long long rows, int mel, float* cnt_scratch, float* dpred) {
  count_span_kernel<<<1, 256, 0, st>>>(span, rows, cnt_scratch);
  F5B_POST_LAUNCH();
  mse_grad_kernel<<<cdiv(rows * mel, 256), 256, 0, st>>>(pred, x1, x0, span, cnt_scratch, mel, rows * mel, dpred);
  F5B_POST_LAUNCH();
}

// ---------------------------------------------------------------------------------------------
// Attention backward (SDPA of modules.py:499-508), fp32, recomputing the probabilities from the saved log-sum-exp:
//   p_ij = exp(s_ij - lse_i), s_ij = scale q_i.k_j;  D_i = dO_i.O_i;  dS_ij = p_ij (dO_i.v_j - D_i) scale
//   dq_i = sum_j dS_ij k_j,  dk_j = sum_i dS_ij q_i,  dv_j = sum_i p_ij dO_i
// One thread owns one query row (lse / dq kernels) or one key row (dk / dv kernels) with its 64-wide vectors in
// registers; the other side streams through shared memory in tiles of 64 rows.  No atomics.
// qkv [batches * rows, 3 * inner] (rotated q, k); d_o / o [batches * rows, inner].  Keys j >= kv_len[b] do not exist.
// ---------------------------------------------------------------------------------------------
static constexpr int AB_T = 64;  // rows per tile / threads per block

__device__ __forceinline__ void load_tile64(float (*dst)[64], const float* __restrict__ src, long long ld, int r0,
                                            int rmax) {
  for (int e = threadIdx.x; e < AB_T * 64; e += AB_T) {
    const int r = e >> 6, c = e & 63;
    dst[r][c] = (r0 + r < rmax) ? src[(long long)(r0 + r) * ld + c] : 0.0f;
  }
}

// lse_i and D_i
__global__ void __launch_bounds__(AB_T)
attn_lse_kernel(const float* __restrict__ qkv, const float* __restrict__ o, const float* __restrict__ d_o, int rows,
                int heads, const int* __restrict__ kv_len, float scale, float* __restrict__ lse, float* __restrict__ D) {
  __shared__ __align__(16) float Ks[AB_T][64];
  const int h = blockIdx.y, b = blockIdx.z, inner = heads * 64;
  const int i = blockIdx.x * AB_T + threadIdx.x;
  const int L = kv_len ? kv_len[b] : rows;
  const float* base = qkv + (long long)b * rows * 3 * inner + h * 64;
  float q[64];
  const bool live = i < rows;
#pragma unroll
  for (int c = 0; c < 64; ++c) q[c] = live ? base[(long long)i * 3 * inner + c] : 0.0f;
  float m = -INFINITY, l = 0.0f;
  for (int j0 = 0; j0 < L; j0 += AB_T) {
    __syncthreads();
    load_tile64(Ks, base + inner, 3 * inner, j0, L);
    __syncthreads();
    const int nj = min(AB_T, L - j0);
    for (int j = 0; j < nj; ++j) {
      float s = 0.0f;
#pragma unroll
      for (int c = 0; c < 64; c += 4) {  // one 16-byte broadcast read per four FMAs
        const float4 kv = *reinterpret_cast<const float4*>(&Ks[j][c]);
        s = fmaf(q[c], kv.x, s); s = fmaf(q[c + 1], kv.y, s); s = fmaf(q[c + 2], kv.z, s); s = fmaf(q[c + 3], kv.w, s);
      }
      s *= scale;
      const float mn = fmaxf(m, s);
      l = l * expf(m - mn) + expf(s - mn);
      m = mn;
    }
  }
  if (!live) return;
  const long long r = (long long)b * rows + i;
  lse[r * heads + h] = m + logf(l);
  float dd = 0.0f;
  const float* op = o + r * inner + h * 64;
  const float* dp = d_o + r * inner + h * 64;
#pragma unroll
  for (int c = 0; c < 64; ++c) dd = fmaf(op[c], dp[c], dd);
  D[r * heads + h] = dd;
}

// dq (written into the q third of dqkv)
__global__ void __launch_bounds__(AB_T)
attn_dq_kernel(const float* __restrict__ qkv, const float* __restrict__ d_o, const float* __restrict__ lse,
               const float* __restrict__ D, int rows, int heads, const int* __restrict__ kv_len, float scale,
               float* __restrict__ dqkv) {
  __shared__ __align__(16) float Ks[AB_T][64], Vs[AB_T][64];
  const int h = blockIdx.y, b = blockIdx.z, inner = heads * 64;
  const int i = blockIdx.x * AB_T + threadIdx.x;
  const int L = kv_len ? kv_len[b] : rows;
  const float* base = qkv + (long long)b * rows * 3 * inner + h * 64;
  const bool live = i < rows;
  const long long r = (long long)b * rows + (live ? i : 0);
  float q[64], g[64], dq[64];
#pragma unroll
  for (int c = 0; c < 64; ++c) {
    q[c] = live ? base[(long long)i * 3 * inner + c] : 0.0f;
    g[c] = live ? d_o[r * inner + h * 64 + c] : 0.0f;
    dq[c] = 0.0f;
  }
  const float ls = live ? lse[r * heads + h] : 0.0f, Di = live ? D[r * heads + h] : 0.0f;
  for (int j0 = 0; j0 < L; j0 += AB_T) {
    __syncthreads();
    load_tile64(Ks, base + inner, 3 * inner, j0, L);
    load_tile64(Vs, base + 2 * inner, 3 * inner, j0, L);
    __syncthreads();
    const int nj = min(AB_T, L - j0);
    for (int j = 0; j < nj; ++j) {
      float s = 0.0f, dp = 0.0f;
#pragma unroll
      for (int c = 0; c < 64; c += 4) {
        const float4 kv = *reinterpret_cast<const float4*>(&Ks[j][c]);
        const float4 vv = *reinterpret_cast<const float4*>(&Vs[j][c]);
        s = fmaf(q[c], kv.x, s); s = fmaf(q[c + 1], kv.y, s); s = fmaf(q[c + 2], kv.z, s); s = fmaf(q[c + 3], kv.w, s);
        dp = fmaf(g[c], vv.x, dp); dp = fmaf(g[c + 1], vv.y, dp); dp = fmaf(g[c + 2], vv.z, dp); dp = fmaf(g[c + 3], vv.w, dp);
      }
      const float ds = expf(s * scale - ls) * (dp - Di) * scale;
#pragma unroll
      for (int c = 0; c < 64; c += 4) {
        const float4 kv = *reinterpret_cast<const float4*>(&Ks[j][c]);
        dq[c] = fmaf(ds, kv.x, dq[c]); dq[c + 1] = fmaf(ds, kv.y, dq[c + 1]);
        dq[c + 2] = fmaf(ds, kv.z, dq[c + 2]); dq[c + 3] = fmaf(ds, kv.w, dq[c + 3]);
      }
    }
  }
  if (!live) return;
  float* out = dqkv + r * 3 * inner + h * 64;
#pragma unroll
  for (int c = 0; c < 64; ++c) out[c] = dq[c];
}

// dk (WHICH = 0) or dv (WHICH = 1) of key row j, streaming the query side
template <int WHICH>
__global__ void __launch_bounds__(AB_T)
attn_dkv_kernel(const float* __restrict__ qkv, const float* __restrict__ d_o, const float* __restrict__ lse,
                const float* __restrict__ D, int rows, int heads, const int* __restrict__ kv_len, float scale,
                float* __restrict__ dqkv) {
  __shared__ __align__(16) float Qs[AB_T][64], Gs[AB_T][64];
  __shared__ float Ls[AB_T], Ds[AB_T];
  const int h = blockIdx.y, b = blockIdx.z, inner = heads * 64;
  const int j = blockIdx.x * AB_T + threadIdx.x;
  const int L = kv_len ? kv_len[b] : rows;
  const float* base = qkv + (long long)b * rows * 3 * inner + h * 64;
  const bool live = j < L;
  float k[64], v[64], acc[64];
#pragma unroll
  for (int c = 0; c < 64; ++c) {
    k[c] = live ? base[(long long)j * 3 * inner + inner + c] : 0.0f;
    v[c] = (WHICH == 0 && live) ? base[(long long)j * 3 * inner + 2 * inner + c] : 0.0f;
    acc[c] = 0.0f;
  }
  for (int i0 = 0; i0 < rows; i0 += AB_T) {
    __syncthreads();
    load_tile64(Qs, base, 3 * inner, i0, rows);
    load_tile64(Gs, d_o + (long long)b * rows * inner + h * 64, inner, i0, rows);
    if (i0 + (int)threadIdx.x < rows) {
      Ls[threadIdx.x] = lse[((long long)b * rows + i0 + threadIdx.x) * heads + h];
      Ds[threadIdx.x] = D[((long long)b * rows + i0 + threadIdx.x) * heads + h];
    }
    __syncthreads();
    const int ni = min(AB_T, rows - i0);
    for (int i = 0; i < ni; ++i) {
      float s = 0.0f;
#pragma unroll
      for (int c = 0; c < 64; c += 4) {
        const float4 qv = *reinterpret_cast<const float4*>(&Qs[i][c]);
        s = fmaf(qv.x, k[c], s); s = fmaf(qv.y, k[c + 1], s); s = fmaf(qv.z, k[c + 2], s); s = fmaf(qv.w, k[c + 3], s);
      }
      const float p = expf(s * scale - Ls[i]);
      if (WHICH == 1) {
#pragma unroll
        for (int c = 0; c < 64; c += 4) {
          const float4 gv = *reinterpret_cast<const float4*>(&Gs[i][c]);
          acc[c] = fmaf(p, gv.x, acc[c]); acc[c + 1] = fmaf(p, gv.y, acc[c + 1]);
          acc[c + 2] = fmaf(p, gv.z, acc[c + 2]); acc[c + 3] = fmaf(p, gv.w, acc[c + 3]);
        }
      } else {
        float dp = 0.0f;
#pragma unroll
        for (int c = 0; c < 64; c += 4) {
          const float4 gv = *reinterpret_cast<const float4*>(&Gs[i][c]);
          dp = fmaf(gv.x, v[c], dp); dp = fmaf(gv.y, v[c + 1], dp); dp = fmaf(gv.z, v[c + 2], dp); dp = fmaf(gv.w, v[c + 3], dp);
        }
        const float ds = p * (dp - Ds[i]) * scale;
#pragma unroll
        for (int c = 0; c < 64; c += 4) {
          const float4 qv = *reinterpret_cast<const float4*>(&Qs[i][c]);
          acc[c] = fmaf(ds, qv.x, acc[c]); acc[c + 1] = fmaf(ds, qv.y, acc[c + 1]);
          acc[c + 2] = fmaf(ds, qv.z, acc[c + 2]); acc[c + 3] = fmaf(ds, qv.w, acc[c + 3]);
        }
      }
    }
  }
  if (j >= rows) return;
  float* out = dqkv + ((long long)b * rows + j) * 3 * inner + (WHICH == 0 ? inner : 2 * inner) + h * 64;
#pragma unroll
  for (int c = 0; c < 64; ++c) out[c] = live ? acc[c] : 0.0f;  // keys beyond kv_len took no part: zero gradient
}

void attention_backward(cudaStream_t st, const float* qkv, const float* o, const float* d_o, int batches, int rows,
                        int heads, const int* kv_len, float scale, float* lse, float* D, float* dqkv) {
  dim3 grid(cdiv(rows, AB_T), heads, batches);
  attn_lse_kernel<<<grid, AB_T, 0, st>>>(qkv, o, d_o, rows, heads, kv_len, scale, lse, D);
  F5B_POST_LAUNCH();
  attn_dq_kernel<<<grid, AB_T, 0, st>>>(qkv, d_o, lse, D, rows, heads, kv_len, scale, dqkv);
  F5B_POST_LAUNCH();
  attn_dkv_kernel<0><<<grid, AB_T, 0, st>>>(qkv, d_o, lse, D, rows, heads, kv_len, scale, dqkv);
  F5B_POST_LAUNCH();
  attn_dkv_kernel<1><<<grid, AB_T, 0, st>>>(qkv, d_o, lse, D, rows, heads, kv_len, scale, dqkv);
  F5B_POST_LAUNCH();
}

}  // namespace f5b

namespace f5b {

// ---------------------------------------------------------------------------------------------
// 16-bit operand preparation for the tensor-core tier of the backward GEMMs (tc_gemm computes A[M,K] W[N,K]^T):
//   cast16           : fp32 [R, C] -> 16-bit [R, ldd], columns >= C zero (ldd = C rounded up to 8: TMA stride rule)
//   transpose_cast16 : fp32 [R, C] -> 16-bit [C, ldd], ldd >= R, columns >= R zero (32 x 32 tiles through shared memory)
// fp16 = 1: IEEE half (forward operands), 0: bf16 (gradient operands need the exponent range)
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ uint16_t to16(float v, int fp16) {
  if (fp16) {
    __half h = __float2half_rn(v);
    return *reinterpret_cast<uint16_t*>(&h);
  }
  bf16 h = __float2bfloat16_rn(v);
  return *reinterpret_cast<uint16_t*>(&h);
}
__global__ void __launch_bounds__(256)
cast16_kernel(const float* __restrict__ src, long long lds, long long R, int C, uint16_t* __restrict__ dst, long long ldd,
              int fp16) {
  const long long i = (long long)blockIdx.x * 256 + threadIdx.x;
  if (i >= R * ldd) return;
  const long long r = i / ldd;
  const int c = (int)(i - r * ldd);
  dst[i] = c < C ? to16(src[r * lds + c], fp16) : (uint16_t)0;
}
// C, lds, ldd multiples of 4 and 16-byte aligned rows: one float4 in, four 16-bit values out per thread
__global__ void __launch_bounds__(256)
cast16x4_kernel(const float* __restrict__ src, long long lds, long long R, int C, uint16_t* __restrict__ dst, long long ldd,
                int fp16) {
  const long long i = (long long)blockIdx.x * 256 + threadIdx.x;
  const long long q = ldd / 4;
  if (i >= R * q) return;
  const long long r = i / q;
  const int c = (int)(i - r * q) * 4;
  uint2 o = make_uint2(0u, 0u);
  if (c < C) {
    const float4 v = *reinterpret_cast<const float4*>(src + r * lds + c);
    o.x = (uint32_t)to16(v.x, fp16) | ((uint32_t)to16(v.y, fp16) << 16);
    o.y = (uint32_t)to16(v.z, fp16) | ((uint32_t)to16(v.w, fp16) << 16);
  }
  *reinterpret_cast<uint2*>(dst + r * ldd + c) = o;
}
void cast16(cudaStream_t st, const float* src, long long lds, long long R, int C, void* dst, long long ldd, int fp16) {
  if (C % 4 == 0 && lds % 4 == 0 && ldd % 4 == 0 && al16(src) && (reinterpret_cast<uintptr_t>(dst) & 7) == 0) {
    cast16x4_kernel<<<cdiv(R * (ldd / 4), 256), 256, 0, st>>>(src, lds, R, C, reinterpret_cast<uint16_t*>(dst), ldd, fp16);
    F5B_POST_LAUNCH();
    return;
  }
  cast16_kernel<<<cdiv(R * ldd, 256), 256, 0, st>>>(src, lds, R, C, reinterpret_cast<uint16_t*>(dst), ldd, fp16);
  F5B_POST_LAUNCH();
}
__global__ void __launch_bounds__(256)
transpose_cast16_kernel(const float* __restrict__ src, long long lds, long long R, int C, uint16_t* __restrict__ dst,
                        long long ldd, int fp16) {
  __shared__ float tile[32][33];
  const long long r0 = (long long)blockIdx.x * 32;
  const int c0 = blockIdx.y * 32;
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;  // 32 x 8
  for (int k = ty; k < 32; k += 8) {
    const long long r = r0 + k;
    const int c = c0 + tx;
    tile[k][tx] = (r < R && c < C) ? src[r * lds + c] : 0.0f;
  }
  __syncthreads();
  for (int k = ty; k < 32; k += 8) {
    const int c = c0 + k;
    const long long r = r0 + tx;
    if (c < C && r < ldd) dst[(long long)c * ldd + r] = to16(tile[tx][k], fp16);  // rows >= R were read as zero
  }
}
// 64 source rows x 32 source columns per CTA: a warp reads 128 B of a source row and writes 128 B (64 packed 16-bit
// values) of a destination row -- the 32 x 32 tile above wrote 64-byte pieces
__global__ void __launch_bounds__(256)
transpose_cast16w_kernel(const float* __restrict__ src, long long lds, long long R, int C, uint16_t* __restrict__ dst,
                         long long ldd, int fp16) {
  __shared__ float tile[32][65];  // [source column][source row]
  const long long r0 = (long long)blockIdx.x * 64;
  const int c0 = blockIdx.y * 32;
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
  for (int k = ty; k < 64; k += 8) {
    const long long r = r0 + k;
    const int c = c0 + tx;
    tile[tx][k] = (r < R && c < C) ? src[r * lds + c] : 0.0f;
  }
  __syncthreads();
  for (int k = ty; k < 32; k += 8) {
    const int c = c0 + k;
    const long long r = r0 + 2 * tx;  // ldd is even (asserted by the launcher): r + 1 < ldd whenever r < ldd
    if (c < C && r < ldd)
      *reinterpret_cast<uint32_t*>(dst + (long long)c * ldd + r) =
          (uint32_t)to16(tile[k][2 * tx], fp16) | ((uint32_t)to16(tile[k][2 * tx + 1], fp16) << 16);
  }
}
__global__ void __launch_bounds__(256)
widen_bf16x4_kernel(const uint2* __restrict__ src, long long n4, float4* __restrict__ dst) {
  const long long i = (long long)blockIdx.x * 256 + threadIdx.x;
  if (i >= n4) return;
  const uint2 v = src[i];
  dst[i] = make_float4(__uint_as_float(v.x << 16), __uint_as_float(v.x & 0xffff0000u), __uint_as_float(v.y << 16),
                       __uint_as_float(v.y & 0xffff0000u));
}
__global__ void __launch_bounds__(256)
widen_bf16_kernel(const bf16* __restrict__ src, long long n, float* __restrict__ dst) {
  const long long i = (long long)blockIdx.x * 256 + threadIdx.x;
  if (i < n) dst[i] = __bfloat162float(src[i]);
}
void widen_bf16(cudaStream_t st, const void* src, long long n, float* dst) {
  if (n % 4 == 0 && al16(dst) && (reinterpret_cast<uintptr_t>(src) & 7) == 0)
    widen_bf16x4_kernel<<<cdiv(n / 4, 256), 256, 0, st>>>(reinterpret_cast<const uint2*>(src), n / 4,
                                                        reinterpret_cast<float4*>(dst));
  else
    widen_bf16_kernel<<<cdiv(n, 256), 256, 0, st>>>(reinterpret_cast<const bf16*>(src), n, dst);
  F5B_POST_LAUNCH();
}
// One pass over an upstream gradient dY [R, C] (fp32) for everything its linear layer's backward needs from it:
//   dst16 [R, C]   bf16 copy, the K-major A operand of the dgrad GEMM
//   dstT  [C, ldT] bf16 transpose (columns >= R zero), the K-major A operand of the wgrad GEMM
//   part  [row tiles, C] column sums of the 64-row tiles -> bias gradient (summed in tile order by colsum_final)
// instead of a cast pass, a transpose pass and a column-sum pass that each read dY.  64 x 64 tiles, C % 64 == 0.
__global__ void __launch_bounds__(256)
prep_dy16_kernel(const float* __restrict__ src, long long R, int C, uint16_t* __restrict__ dst16,
                 uint16_t* __restrict__ dstT, long long ldT, float* __restrict__ part) {
  __shared__ float tile[64][65];  // [column][row]
  const long long r0 = (long long)blockIdx.x * 64;
  const int c0 = blockIdx.y * 64;
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
  for (int k = ty; k < 64; k += 8) {
    const long long r = r0 + k;
    float2 v = make_float2(0.f, 0.f);
    if (r < R) {
      v = *reinterpret_cast<const float2*>(src + r * C + c0 + 2 * tx);
      *reinterpret_cast<uint32_t*>(dst16 + r * C + c0 + 2 * tx) = (uint32_t)to16(v.x, 0) | ((uint32_t)to16(v.y, 0) << 16);
    }
    tile[2 * tx][k] = v.x;
    tile[2 * tx + 1][k] = v.y;
  }
  __syncthreads();
  for (int k = ty; k < 64; k += 8) {
    const int c = c0 + k;
    const float a = tile[k][2 * tx], b = tile[k][2 * tx + 1];
    const long long r = r0 + 2 * tx;
    if (r < ldT) *reinterpret_cast<uint32_t*>(dstT + (long long)c * ldT + r) = (uint32_t)to16(a, 0) | ((uint32_t)to16(b, 0) << 16);
    float s = a + b;
#pragma unroll
    for (int off = 16; off > 0; off >>= 1) s += __shfl_xor_sync(0xffffffffu, s, off);
    if (tx == 0) part[(long long)blockIdx.x * C + c] = s;
  }
}
// returns false (nothing done) when the shape does not fit the fused form
bool prep_dy16(cudaStream_t st, const float* src, long long R, int C, void* dst16, void* dstT, long long ldT, float* part,
               size_t part_floats, float* colsum_out) {
  const long long tiles = (ldT + 63) / 64;
  if (C % 64 != 0 || ldT % 2 != 0 || (size_t)tiles * C > part_floats || part == nullptr) return false;
  prep_dy16_kernel<<<dim3((unsigned)tiles, C / 64), 256, 0, st>>>(src, R, C, reinterpret_cast<uint16_t*>(dst16),
                                                                 reinterpret_cast<uint16_t*>(dstT), ldT, part);
  F5B_POST_LAUNCH();
  // the tile sums ([tiles, C], 600 x 3072 at C5 sizes) are added with the 16-byte, four-row-group kernel: the
  // one-thread-per-column loop of colsum_final took 35 us here
  colsum_seg4_kernel<<<dim3(cdiv(C, 256), 1), 256, 0, st>>>(part, nullptr, (int)tiles, (int)tiles, C, colsum_out, 0);
  F5B_POST_LAUNCH();
  return true;
}
void transpose_cast16(cudaStream_t st, const float* src, long long lds, long long R, int C, void* dst, long long ldd,
                      int fp16) {
  if (ldd % 2 == 0 && (reinterpret_cast<uintptr_t>(dst) & 3) == 0) {
    transpose_cast16w_kernel<<<dim3(cdiv(ldd, 64), cdiv(C, 32)), 256, 0, st>>>(src, lds, R, C,
                                                                              reinterpret_cast<uint16_t*>(dst), ldd, fp16);
    F5B_POST_LAUNCH();
    return;
  }
  transpose_cast16_kernel<<<dim3(cdiv(ldd, 32), cdiv(C, 32)), 256, 0, st>>>(src, lds, R, C, reinterpret_cast<uint16_t*>(dst),
                                                                           ldd, fp16);
  F5B_POST_LAUNCH();
}

}  // namespace f5b
