// Backward kernels of the embedding stage of the training step: InputEmbedding (dit.py:121-140: Linear over
// cat(x, cond, text) + ConvPositionEmbedding, modules.py:170-196) and TextEmbedding (dit.py:32-115: token lookup, sinus
// positions, ConvNeXt-V2 blocks with GRN, modules.py:231-275).  fp32 CUDA-core arithmetic, deterministic (every
// reduction has one owner thread or a fixed tree).  Activations are channels-last: [batch, rows, channels].
#include "train.h"

#include <algorithm>

namespace f5b {

static inline unsigned cdiv2(long long a, long long b) { return (unsigned)((a + b - 1) / b); }

// ---------------------------------------------------------------------------------------------
// Mish (modules.py:176,178) and exact GELU (modules.py:264)
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
mish_fwd_mask_kernel(const float* __restrict__ z, float* __restrict__ y, int rows, int cols, const int* __restrict__ row_len,
                     long long total) {
  const long long i = (long long)blockIdx.x * 256 + threadIdx.x;
  if (i >= total) return;
  const long long m = i / cols;
  const long long b = m / rows;
  const bool dead = row_len != nullptr && (int)(m - b * rows) >= row_len[b];
  y[i] = dead ? 0.0f : act_apply(ACT_MISH, z[i]);
}
// dz = dy * mish'(z), zero on masked rows;  mish'(z) = tanh(sp) + z sigmoid(z) (1 - tanh(sp)^2), sp = softplus(z)
__global__ void __launch_bounds__(256)
mish_bwd_mask_kernel(const float* __restrict__ z, const float* __restrict__ dy, float* __restrict__ dz, int rows, int cols,
                     const int* __restrict__ row_len, long long total) {
  const long long i = (long long)blockIdx.x * 256 + threadIdx.x;
  if (i >= total) return;
  const long long m = i / cols;
  const long long b = m / rows;
  if (row_len != nullptr && (int)(m - b * rows) >= row_len[b]) {
    dz[i] = 0.0f;
    return;
  }
  const float x = z[i];
  const float sp = x > 20.0f ? x : log1pf(expf(x));
  const float t = tanhf(sp), sg = 1.0f / (1.0f + expf(-x));
  dz[i] = dy[i] * (t + x * sg * (1.0f - t * t));
}
void mish_fwd_mask(cudaStream_t st, const float* z, float* y, int batches, int rows, int cols, const int* row_len) {
  const long long total = (long long)batches * rows * cols;
  mish_fwd_mask_kernel<<<cdiv2(total, 256), 256, 0, st>>>(z, y, rows, cols, row_len, total);
  F5B_POST_LAUNCH();
}
void mish_bwd_mask(cudaStream_t st, const float* z, const float* dy, float* dz, int batches, int rows, int cols,
                   const int* row_len) {
  const long long total = (long long)batches * rows * cols;
  mish_bwd_mask_kernel<<<cdiv2(total, 256), 256, 0, st>>>(z, dy, dz, rows, cols, row_len, total);
  F5B_POST_LAUNCH();
}

__global__ void __launch_bounds__(256)
gelu_erf_fwd_kernel(const float* __restrict__ u, float* __restrict__ a, long long n) {
  const long long i = (long long)blockIdx.x * 256 + threadIdx.x;
  if (i < n) a[i] = act_apply(ACT_GELU_ERF, u[i]);
}
__global__ void __launch_bounds__(256)
gelu_erf_bwd_kernel(const float* __restrict__ u, const float* __restrict__ da, float* __restrict__ du, long long n) {
  const long long i = (long long)blockIdx.x * 256 + threadIdx.x;
  if (i >= n) return;
  const float x = u[i];
  du[i] = da[i] * (0.5f * (1.0f + erff(x * 0.7071067811865476f)) + x * 0.3989422804014327f * expf(-0.5f * x * x));
}
void gelu_erf_fwd(cudaStream_t st, const float* u, float* a, long long n) {
  gelu_erf_fwd_kernel<<<cdiv2(n, 256), 256, 0, st>>>(u, a, n);
  F5B_POST_LAUNCH();
}
void gelu_erf_bwd(cudaStream_t st, const float* u, const float* da, float* du, long long n) {
  gelu_erf_bwd_kernel<<<cdiv2(n, 256), 256, 0, st>>>(u, da, du, n);
  F5B_POST_LAUNCH();
}

// ---------------------------------------------------------------------------------------------
// Grouped Conv1d, kernel `taps` (odd), `groups`, zero padding; rows >= row_len[b] of the INPUT read as zero.
//   w [channels(co), cg(ci local), taps] (PyTorch layout);  z[b,t,co] = bias[co] + sum_{tap,ci} x[b,t+tap-pad,ci] w[co,ci_l,tap]
// forward: one thread per (row, co);  dgrad: one thread per (row, ci);  wgrad: one thread per weight, rows in order.
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
gconv_fwd_kernel(const float* __restrict__ x, const float* __restrict__ w, const float* __restrict__ bias, int rows,
                 int C, int cg, int taps, const int* __restrict__ row_len, long long total, float* __restrict__ z) {
  const long long i = (long long)blockIdx.x * 256 + threadIdx.x;
  if (i >= total) return;
  const long long m = i / C;
  const int co = (int)(i - m * C), b = (int)(m / rows), t = (int)(m - (long long)b * rows);
  const int len = row_len ? row_len[b] : rows, pad = taps / 2, c0 = (co / cg) * cg;
  float acc = bias[co];
  for (int tap = 0; tap < taps; ++tap) {
    const int tt = t + tap - pad;
    if (tt < 0 || tt >= len) continue;
    const float* xr = x + ((long long)b * rows + tt) * C + c0;
    const float* wr = w + (long long)co * cg * taps + tap;
    for (int ci = 0; ci < cg; ++ci) acc = fmaf(xr[ci], wr[(long long)ci * taps], acc);
  }
  z[i] = acc;
}
// Depthwise form (groups == C, the ConvNeXt blocks' k = 7): a thread owns one channel of a 64-row slice of one batch
// element and slides a register window down the rows -- every input is read once, coalesced over the channels, no
// index division per element.  DIR = +1: forward (z = bias + sum_k x[t + k - pad] w[k], input rows >= len read as zero);
// DIR = -1: data gradient (dx = sum_k dz[t - k + pad] w[k] for t < len, 0 beyond).
template <int TAPS, int DIR>
__global__ void __launch_bounds__(256)
dwconv_slide_kernel(const float* __restrict__ in, const float* __restrict__ w, const float* __restrict__ bias, int rows,
                    int C, const int* __restrict__ row_len, float* __restrict__ out) {
  constexpr int PAD = TAPS / 2, SLICE = 64;
  const int c = blockIdx.x * 256 + threadIdx.x, t0 = blockIdx.y * SLICE, b = blockIdx.z;
  if (c >= C) return;
  const int len = row_len ? row_len[b] : rows;
  const int in_len = DIR > 0 ? len : rows;  // rows of the input that exist (dz is defined on every row)
  float wk[TAPS], win[TAPS];
#pragma unroll
  for (int k = 0; k < TAPS; ++k) wk[k] = w[(long long)c * TAPS + (DIR > 0 ? k : TAPS - 1 - k)];
  const float* ib = in + (long long)b * rows * C + c;
  float* ob = out + (long long)b * rows * C + c;
  auto ld = [&](int t) { return (t >= 0 && t < in_len) ? ib[(long long)t * C] : 0.0f; };
#pragma unroll
  for (int k = 0; k < TAPS - 1; ++k) win[k + 1] = ld(t0 + k - PAD);  // window holds rows t - PAD .. t + PAD
  const float bv = (DIR > 0 && bias) ? bias[c] : 0.0f;
  const int t1 = min(rows, t0 + SLICE);
  for (int t = t0; t < t1; ++t) {
#pragma unroll
    for (int k = 0; k < TAPS - 1; ++k) win[k] = win[k + 1];
    win[TAPS - 1] = ld(t + PAD);
    float acc = bv;
#pragma unroll
    for (int k = 0; k < TAPS; ++k) acc = fmaf(win[k], wk[k], acc);
    ob[(long long)t * C] = (DIR < 0 && t >= len) ? 0.0f : acc;
  }
}

void gconv_fwd(cudaStream_t st, const float* x, const float* w, const float* bias, int batches, int rows, int C, int groups,
               int taps, const int* row_len, float* z) {
  if (groups == C && taps == 7) {
    dwconv_slide_kernel<7, 1><<<dim3(cdiv2(C, 256), cdiv2(rows, 64), batches), 256, 0, st>>>(x, w, bias, rows, C, row_len, z);
    F5B_POST_LAUNCH();
    return;
  }
  const long long total = (long long)batches * rows * C;
  gconv_fwd_kernel<<<cdiv2(total, 256), 256, 0, st>>>(x, w, bias, rows, C, C / groups, taps, row_len, total, z);
  F5B_POST_LAUNCH();
}

// dx[b,t,ci] = sum_{tap, co in group} dz[b, t - tap + pad, co] w[co, ci_l, tap]   (dz rows >= row_len are zero already)
__global__ void __launch_bounds__(256)
gconv_dgrad_kernel(const float* __restrict__ dz, const float* __restrict__ w, int rows, int C, int cg, int taps,
                   const int* __restrict__ row_len, long long total, float* __restrict__ dx) {
  const long long i = (long long)blockIdx.x * 256 + threadIdx.x;
  if (i >= total) return;
  const long long m = i / C;
  const int ci = (int)(i - m * C), b = (int)(m / rows), t = (int)(m - (long long)b * rows);
  const int len = row_len ? row_len[b] : rows, pad = taps / 2, c0 = (ci / cg) * cg, cl = ci - c0;
  float acc = 0.0f;
  if (t < len) {  // input rows beyond the utterance were read as zero: no gradient
    for (int tap = 0; tap < taps; ++tap) {
      const int tt = t - tap + pad;
      if (tt < 0 || tt >= rows) continue;
      const float* gr = dz + ((long long)b * rows + tt) * C + c0;
      const float* wr = w + ((long long)c0 * cg + cl) * taps + tap;
      for (int co = 0; co < cg; ++co) acc = fmaf(gr[co], wr[(long long)co * cg * taps], acc);
    }
  }
  dx[i] = acc;
}
void gconv_dgrad(cudaStream_t st, const float* dz, const float* w, int batches, int rows, int C, int groups, int taps,
                 const int* row_len, float* dx) {
  if (groups == C && taps == 7) {
    dwconv_slide_kernel<7, -1><<<dim3(cdiv2(C, 256), cdiv2(rows, 64), batches), 256, 0, st>>>(dz, w, nullptr, rows, C, row_len, dx);
    F5B_POST_LAUNCH();
    return;
  }
  const long long total = (long long)batches * rows * C;
  gconv_dgrad_kernel<<<cdiv2(total, 256), 256, 0, st>>>(dz, w, rows, C, C / groups, taps, row_len, total, dx);
  F5B_POST_LAUNCH();
}

// dw[co, ci_l, tap] = sum_{b,t} dz[b,t,co] x[b, t + tap - pad, ci]
__global__ void __launch_bounds__(256)
gconv_wgrad_kernel(const float* __restrict__ dz, const float* __restrict__ x, int batches, int rows, int C, int cg,
                   int taps, const int* __restrict__ row_len, long long total, float* __restrict__ dw) {
  const long long i = (long long)blockIdx.x * 256 + threadIdx.x;
  if (i >= total) return;
  const int tap = (int)(i % taps);
  const long long r = i / taps;
  const int cl = (int)(r % cg), co = (int)(r / cg), ci = (co / cg) * cg + cl, pad = taps / 2;
  float acc = 0.0f;
  for (int b = 0; b < batches; ++b) {
    const int len = row_len ? row_len[b] : rows;
    for (int t = 0; t < rows; ++t) {
      const int tt = t + tap - pad;
      if (tt < 0 || tt >= len) continue;
      acc = fmaf(dz[((long long)b * rows + t) * C + co], x[((long long)b * rows + tt) * C + ci], acc);
    }
  }
  dw[i] = acc;
}
// 64 channels per group (dim 1024): one CTA per (group, tap) accumulates its 64 x 64 block of dw over all rows through
// shared-memory tiles of 16 rows -- 4 x 4 outputs per thread instead of one thread per weight walking every row
__global__ void __launch_bounds__(256)
gconv_wgrad64_kernel(const float* __restrict__ dz, const float* __restrict__ x, int batches, int rows, int C, int taps,
                     const int* __restrict__ row_len, float* __restrict__ dw) {
  // 32-row tiles, rows of 64 floats (16-byte aligned): each thread reads its 4 + 4 operands of a row as two 128-bit
  // shared loads (the 65-float padded rows of the first version made them eight scalar loads per 16 FMAs)
  __shared__ __align__(16) float Gs[32][64], Xs[32][64];
  const int g = blockIdx.x, tap = blockIdx.y, pad = taps / 2;
  const int tid = threadIdx.x, tx = tid & 15, ty = tid >> 4;
  float acc[4][4] = {};
  for (int b = 0; b < batches; ++b) {
    const int len = row_len ? row_len[b] : rows;
    for (int t0 = 0; t0 < rows; t0 += 32) {
      for (int e = tid; e < 32 * 16; e += 256) {  // one float4 per thread and pass
        const int r = e >> 4, c = (e & 15) * 4, t = t0 + r, tt = t + tap - pad;
        float4 gv = make_float4(0.f, 0.f, 0.f, 0.f), xv = gv;
        if (t < rows) {
          gv = *reinterpret_cast<const float4*>(dz + ((long long)b * rows + t) * C + g * 64 + c);
          if (tt >= 0 && tt < len) xv = *reinterpret_cast<const float4*>(x + ((long long)b * rows + tt) * C + g * 64 + c);
        }
        *reinterpret_cast<float4*>(&Gs[r][c]) = gv;
        *reinterpret_cast<float4*>(&Xs[r][c]) = xv;
      }
      __syncthreads();
#pragma unroll 8
      for (int r = 0; r < 32; ++r) {
        const float4 av = *reinterpret_cast<const float4*>(&Gs[r][ty * 4]);
        const float4 vv = *reinterpret_cast<const float4*>(&Xs[r][tx * 4]);
        const float a[4] = {av.x, av.y, av.z, av.w}, v[4] = {vv.x, vv.y, vv.z, vv.w};
#pragma unroll
        for (int i = 0; i < 4; ++i)
#pragma unroll
          for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(a[i], v[j], acc[i][j]);
      }
      __syncthreads();
    }
  }
#pragma unroll
  for (int i = 0; i < 4; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j)
      dw[((long long)(g * 64 + ty * 4 + i) * 64 + tx * 4 + j) * taps + tap] = acc[i][j];
}
// Depthwise (one channel per group; the ConvNeXt blocks' k = 7 convolution, modules.py:246): the rows are cut into
// `chunks` slices, a thread owns one channel of one slice and keeps its `taps` sums in registers (coalesced over the
// channels), the slices are added in order by a second kernel.  The generic kernel above walked all rows in ONE thread
// per weight: 20 ms per call at 38 400 rows.
template <int TAPS>
__global__ void __launch_bounds__(256)
dwconv_wgrad_part_kernel(const float* __restrict__ dz, const float* __restrict__ x, int rows, int C,
                         const int* __restrict__ row_len, long long M, long long rows_per_chunk, float* __restrict__ part) {
  const int c = blockIdx.x * 256 + threadIdx.x;
  if (c >= C) return;
  const long long m0 = (long long)blockIdx.y * rows_per_chunk, m1 = min(M, m0 + rows_per_chunk);
  constexpr int pad = TAPS / 2;
  float acc[TAPS];
#pragma unroll
  for (int k = 0; k < TAPS; ++k) acc[k] = 0.0f;
  for (long long m = m0; m < m1; ++m) {
    const int b = (int)(m / rows), t = (int)(m - (long long)b * rows);
    const int len = row_len ? row_len[b] : rows;
    const float g = dz[m * C + c];
#pragma unroll
    for (int k = 0; k < TAPS; ++k) {
      const int tt = t + k - pad;
      if (tt >= 0 && tt < len) acc[k] = fmaf(g, x[((long long)b * rows + tt) * C + c], acc[k]);
    }
  }
#pragma unroll
  for (int k = 0; k < TAPS; ++k) part[((long long)blockIdx.y * TAPS + k) * C + c] = acc[k];
}
__global__ void __launch_bounds__(256)
dwconv_wgrad_sum_kernel(const float* __restrict__ part, int chunks, int C, int taps, float* __restrict__ dw) {
  const int i = blockIdx.x * 256 + threadIdx.x;  // = c * taps + tap (PyTorch layout [C, 1, taps])
  if (i >= C * taps) return;
  const int c = i / taps, k = i - c * taps;
  float acc = 0.0f;
  for (int s = 0; s < chunks; ++s) acc += part[((long long)s * taps + k) * C + c];
  dw[i] = acc;
}

void gconv_wgrad(cudaStream_t st, const float* dz, const float* x, int batches, int rows, int C, int groups, int taps,
                 const int* row_len, float* dw) {
  size_t scratch_floats = 0;
  float* scratch = colsum_scratch(&scratch_floats);
  if (groups == C && taps == 7 && scratch != nullptr && scratch_floats >= (size_t)8 * taps * C) {
    const long long M = (long long)batches * rows;
    int chunks = (int)std::min<size_t>(128, scratch_floats / ((size_t)taps * C));
    const long long rpc = (M + chunks - 1) / chunks;
    chunks = (int)((M + rpc - 1) / rpc);
    dwconv_wgrad_part_kernel<7><<<dim3(cdiv2(C, 256), chunks), 256, 0, st>>>(dz, x, rows, C, row_len, M, rpc, scratch);
    F5B_POST_LAUNCH();
    dwconv_wgrad_sum_kernel<<<cdiv2((long long)C * taps, 256), 256, 0, st>>>(scratch, chunks, C, taps, dw);
    F5B_POST_LAUNCH();
    return;
  }
  if (C / groups == 64) {
    gconv_wgrad64_kernel<<<dim3(groups, taps), 256, 0, st>>>(dz, x, batches, rows, C, taps, row_len, dw);
    F5B_POST_LAUNCH();
    return;
  }
  const long long total = (long long)C * (C / groups) * taps;
  gconv_wgrad_kernel<<<cdiv2(total, 256), 256, 0, st>>>(dz, x, batches, rows, C, C / groups, taps, row_len, total, dw);
  F5B_POST_LAUNCH();
}

// ---------------------------------------------------------------------------------------------
// GRN (modules.py:231-240) over the first row_len[b] rows of each batch element:
//   Gx[b,c] = ||a[b,:,c]||_2,  Nx = Gx / (mean_c Gx + 1e-6),  g = gamma (a Nx) + beta + a
// stats [batches, 3, C]: Gx | Nx | (backward) dGx;  mu [batches]
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
grn_gx_kernel(const float* __restrict__ a, int rows, int C, const int* __restrict__ row_len, float* __restrict__ stats) {
  const int c = blockIdx.x * 256 + threadIdx.x, b = blockIdx.y;
  if (c >= C) return;
  const int len = row_len ? row_len[b] : rows;
  float s = 0.0f;
  for (int r = 0; r < len; ++r) {
    const float v = a[((long long)b * rows + r) * C + c];
    s = fmaf(v, v, s);
  }
  stats[(long long)b * 3 * C + c] = sqrtf(s);
}
__global__ void __launch_bounds__(256) grn_nx_kernel2(float* __restrict__ stats, int C, float* __restrict__ mu) {
  __shared__ float red[256];
  float* g = stats + (long long)blockIdx.x * 3 * C;
  float s = 0.0f;
  for (int c = threadIdx.x; c < C; c += 256) s += g[c];
  red[threadIdx.x] = s;
  __syncthreads();
  for (int o = 128; o > 0; o >>= 1) {
    if ((int)threadIdx.x < o) red[threadIdx.x] += red[threadIdx.x + o];
    __syncthreads();
  }
  const float denom = red[0] / C + 1e-6f;
  if (threadIdx.x == 0) mu[blockIdx.x] = denom;
  for (int c = threadIdx.x; c < C; c += 256) g[C + c] = g[c] / denom;
}
__global__ void __launch_bounds__(256)
grn_apply_fwd_kernel(const float* __restrict__ a, const float* __restrict__ stats, const float* __restrict__ gamma,
                     const float* __restrict__ beta, int rows, int C, const int* __restrict__ row_len, long long total,
                     float* __restrict__ g) {
  const long long i = (long long)blockIdx.x * 256 + threadIdx.x;
  if (i >= total) return;
  const long long m = i / C;
  const int c = (int)(i - m * C), b = (int)(m / rows);
  const bool live = !row_len || (int)(m - (long long)b * rows) < row_len[b];
  g[i] = live ? gamma[c] * (a[i] * stats[(long long)b * 3 * C + C + c]) + beta[c] + a[i] : 0.0f;
}
void grn_fwd(cudaStream_t st, const float* a, int batches, int rows, int C, const int* row_len, const float* gamma,
             const float* beta, float* stats, float* mu, float* g) {
  grn_gx_kernel<<<dim3(cdiv2(C, 256), batches), 256, 0, st>>>(a, rows, C, row_len, stats);
  F5B_POST_LAUNCH();
  grn_nx_kernel2<<<batches, 256, 0, st>>>(stats, C, mu);
  F5B_POST_LAUNCH();
  const long long total = (long long)batches * rows * C;
  grn_apply_fwd_kernel<<<cdiv2(total, 256), 256, 0, st>>>(a, stats, gamma, beta, rows, C, row_len, total, g);
  F5B_POST_LAUNCH();
}

// per (b, c): S1 = sum_r dg a, S0 = sum_r dg;  part [batches, 3, C]: dNx = gamma S1 | dgamma part = S1 Nx | dbeta part = S0
__global__ void __launch_bounds__(256)
grn_bwd_sums_kernel(const float* __restrict__ dg, const float* __restrict__ a, const float* __restrict__ stats,
                    const float* __restrict__ gamma, int rows, int C, const int* __restrict__ row_len,
                    float* __restrict__ part) {
  const int c = blockIdx.x * 256 + threadIdx.x, b = blockIdx.y;
  if (c >= C) return;
  const int len = row_len ? row_len[b] : rows;
  float s1 = 0.0f, s0 = 0.0f;
  for (int r = 0; r < len; ++r) {
    const long long i = ((long long)b * rows + r) * C + c;
    s1 = fmaf(dg[i], a[i], s1);
    s0 += dg[i];
  }
  float* p = part + (long long)b * 3 * C;
  p[c] = gamma[c] * s1;
  p[C + c] = s1 * stats[(long long)b * 3 * C + C + c];
  p[2 * C + c] = s0;
}
// per b: T = sum_c dNx Gx;  dGx[c] = dNx[c] / mu - T / (mu^2 C)   (mu already includes the 1e-6)
__global__ void __launch_bounds__(256)
grn_bwd_dgx_kernel(const float* __restrict__ part, float* __restrict__ stats, const float* __restrict__ mu, int C) {
  __shared__ float red[256];
  const int b = blockIdx.x;
  const float* dnx = part + (long long)b * 3 * C;
  float* sb = stats + (long long)b * 3 * C;
  float s = 0.0f;
  for (int c = threadIdx.x; c < C; c += 256) s = fmaf(dnx[c], sb[c], s);
  red[threadIdx.x] = s;
  __syncthreads();
  for (int o = 128; o > 0; o >>= 1) {
    if ((int)threadIdx.x < o) red[threadIdx.x] += red[threadIdx.x + o];
    __syncthreads();
  }
  const float m = mu[b], T = red[0];
  for (int c = threadIdx.x; c < C; c += 256) sb[2 * C + c] = dnx[c] / m - T / (m * m * C);
}
__global__ void __launch_bounds__(256)
grn_bwd_apply_kernel(const float* __restrict__ dg, const float* __restrict__ a, const float* __restrict__ stats,
                     const float* __restrict__ gamma, int rows, int C, const int* __restrict__ row_len, long long total,
                     float* __restrict__ da) {
  const long long i = (long long)blockIdx.x * 256 + threadIdx.x;
  if (i >= total) return;
  const long long m = i / C;
  const int c = (int)(i - m * C), b = (int)(m / rows);
  if (row_len && (int)(m - (long long)b * rows) >= row_len[b]) {
    da[i] = 0.0f;
    return;
  }
  const float* sb = stats + (long long)b * 3 * C;
  const float gx = sb[c];
  da[i] = dg[i] * (1.0f + gamma[c] * sb[C + c]) + (gx > 0.0f ? sb[2 * C + c] * a[i] / gx : 0.0f);
}
// dgamma[c] = sum_b part[b,1,c], dbeta[c] = sum_b part[b,2,c]
__global__ void __launch_bounds__(256)
grn_bwd_params_kernel(const float* __restrict__ part, int batches, int C, float* __restrict__ dgamma,
                      float* __restrict__ dbeta) {
  const int c = blockIdx.x * 256 + threadIdx.x;
  if (c >= C) return;
  float g = 0.0f, bt = 0.0f;
  for (int b = 0; b < batches; ++b) {
    g += part[(long long)b * 3 * C + C + c];
    bt += part[(long long)b * 3 * C + 2 * C + c];
  }
  dgamma[c] = g;
  dbeta[c] = bt;
}
void grn_bwd(cudaStream_t st, const float* dg, const float* a, int batches, int rows, int C, const int* row_len,
             const float* gamma, float* stats, const float* mu, float* part, float* da, float* dgamma, float* dbeta) {
  grn_bwd_sums_kernel<<<dim3(cdiv2(C, 256), batches), 256, 0, st>>>(dg, a, stats, gamma, rows, C, row_len, part);
  F5B_POST_LAUNCH();
  grn_bwd_dgx_kernel<<<batches, 256, 0, st>>>(part, stats, mu, C);
  F5B_POST_LAUNCH();
  const long long total = (long long)batches * rows * C;
  grn_bwd_apply_kernel<<<cdiv2(total, 256), 256, 0, st>>>(dg, a, stats, gamma, rows, C, row_len, total, da);
  F5B_POST_LAUNCH();
  grn_bwd_params_kernel<<<cdiv2(C, 256), 256, 0, st>>>(part, batches, C, dgamma, dbeta);
  F5B_POST_LAUNCH();
}

// ---------------------------------------------------------------------------------------------
// Token embedding backward (dit.py:100: nn.Embedding after the +1 shift; filler id 0):
//   dtable[v, c] = sum over the rows whose (effective) token is v and whose value was not masked of dte[row, c]
// The row's token is recomputed exactly as text_lookup_kernel (rowwise.cu) derives it.
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
embed_bwd_kernel(const float* __restrict__ dte, const long long* __restrict__ text, int B, int nt, int drop_text,
                 int rows, int td, const int* __restrict__ row_len, int mask_padding, int masked_fill_applies,
                 int table_rows, float* __restrict__ dtable) {
  const int i = blockIdx.x * 256 + threadIdx.x;
  if (i >= table_rows * td) return;
  const int v = i / td, c = i - v * td;
  float acc = 0.0f;
  // rows at or beyond nt carry the filler token 0 and count as padding: they reach row 0 of the table only, and only
  // when padding is not masked out -- every other (v, c) stops at the text length (170 instead of ~1000 rows)
  const bool filler_counts = v == 0 && !(masked_fill_applies && mask_padding);
  for (int b = 0; b < B; ++b) {
    int len = row_len ? row_len[b] : rows;
    if (!filler_counts && len > nt) len = nt;
    for (int pos = 0; pos < len; ++pos) {
      long long tok = pos < nt ? text[(long long)b * nt + pos] + 1 : 0;
      if (tok < 0 || tok >= table_rows) tok = 0;
      const bool is_pad = tok == 0;
      if (drop_text) tok = 0;
      if (tok != v || (masked_fill_applies && mask_padding && is_pad)) continue;
      acc += dte[((long long)b * rows + pos) * td + c];
    }
  }
  dtable[i] = acc;
}
void embed_bwd(cudaStream_t st, const float* dte, const long long* text, int B, int nt, int drop_text, int rows, int td,
               const int* row_len, int mask_padding, int masked_fill_applies, int table_rows, float* dtable) {
  embed_bwd_kernel<<<cdiv2((long long)table_rows * td, 256), 256, 0, st>>>(dte, text, B, nt, drop_text, rows, td, row_len,
                                                                          mask_padding, masked_fill_applies, table_rows, dtable);
  F5B_POST_LAUNCH();
}

// Weights of the data-gradient of a grouped Conv1d as a forward convolution (tensor-core tier, implicit-GEMM layout of
// tc_gemm's conv mode: [c_out][tap * cg + c_in]):  dx = conv(dz, w'),  w'[ci][tap' * cg + co_l] = w[g cg + co_l][ci_l][taps-1-tap']
__global__ void conv_w_flipT_bf16_kernel(const float* __restrict__ w, bf16* __restrict__ out, int C, int cg, int taps) {
  const long long i = (long long)blockIdx.x * 256 + threadIdx.x;
  if (i >= (long long)C * cg * taps) return;
  const int col = (int)(i % (cg * taps)), ci = (int)(i / (cg * taps));
  const int tp = col / cg, co_l = col - tp * cg, g = ci / cg, ci_l = ci - g * cg;
  out[i] = __float2bfloat16_rn(w[((long long)(g * cg + co_l) * cg + ci_l) * taps + (taps - 1 - tp)]);
}
void conv_w_flipT_bf16(cudaStream_t st, const float* w, void* out, int C, int groups, int taps) {
  const int cg = C / groups;
  conv_w_flipT_bf16_kernel<<<cdiv2((long long)C * cg * taps, 256), 256, 0, st>>>(w, reinterpret_cast<bf16*>(out), C, cg, taps);
  F5B_POST_LAUNCH();
}

// out[c] = v[c] + k
__global__ void vec_add_const_kernel(const float* __restrict__ v, float k, int n, float* __restrict__ out) {
  const int i = blockIdx.x * 256 + threadIdx.x;
  if (i < n) out[i] = v[i] + k;
}
void vec_add_const(cudaStream_t st, const float* v, float k, int n, float* out) {
  vec_add_const_kernel<<<cdiv2(n, 256), 256, 0, st>>>(v, k, n, out);
  F5B_POST_LAUNCH();
}

// y[m, 0:w) = x[m, off : off + w)  (strided column slice) and the reverse scatter
__global__ void __launch_bounds__(256)
col_slice_kernel(const float* __restrict__ x, long long ldx, int off, int w, long long rows, float* __restrict__ y,
                 long long ldy, int yoff) {
  const long long i = (long long)blockIdx.x * 256 + threadIdx.x;
  if (i >= rows * w) return;
  const long long m = i / w;
  const int c = (int)(i - m * w);
  y[m * ldy + yoff + c] = x[m * ldx + off + c];
}
void col_slice(cudaStream_t st, const float* x, long long ldx, int off, int w, long long rows, float* y, long long ldy,
               int yoff) {
  col_slice_kernel<<<cdiv2(rows * w, 256), 256, 0, st>>>(x, ldx, off, w, rows, y, ldy, yoff);
  F5B_POST_LAUNCH();
}

}  // namespace f5b
