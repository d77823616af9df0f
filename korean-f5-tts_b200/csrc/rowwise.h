// Memory-bound kernels of the CFM.sample path (declarations).  T = activation type handed to the next GEMM
// (float in fp32 mode, bf16 in bf16 mode); the residual stream, ODE state and all statistics are fp32.
#pragma once
#include "common.cuh"

namespace f5b {

// Output tag for the precise bf16 contraction of proj_out: a row of width d is written as three bf16 segments
// [hi | lo | hi] (hi = bf16(v), lo = bf16(v - hi)), to be contracted with weights laid out [W_hi | W_hi | W_lo]:
// sum = hi*W_hi + lo*W_hi + hi*W_lo, i.e. ~16 mantissa bits on both operands at 3x the (tiny) K.
struct bf16x3 { bf16 raw; };

// out[m, :] = LN(x[m, :], eps) * (mul_off + mul[b, :]) + add[b, :]      (reference modules.py:307-342, 256-259)
//   b = (m / rows_per_batch) % bmod; mul/add may be null (-> 1 / 0).
template <typename T>
void ln_mod(cudaStream_t st, const float* x, T* out, long long M, int d, const float* mul, const float* add,
            float mul_off, long long vec_bstride, int rows_per_batch, int bmod, float eps);

// x_transformers RMSNorm: out = x / max(||x||, 1e-12) * sqrt(d) * g     (reference unett.py:154,168,185)
template <typename T>
void xrms_norm(cudaStream_t st, const float* x, T* out, long long M, int d, const float* g);

// ConvNeXt front half: depthwise Conv1d(k=7, pad=3) along rows of each batch element, + bias, LayerNorm(affine).
//   x fp32 [batches, rows, d]; rows >= row_len[b] read as zero; w_t [7][d].   (modules.py:256-259; vocos ConvNeXtBlock)
template <typename T>
void dwconv7_ln(cudaStream_t st, const float* x, T* out, int batches, int rows, int d, const int* row_len,
                const float* w_t, const float* bias, const float* ln_w, const float* ln_b, float eps);

// GRN (modules.py:231-240) in place on h [batches, rows, C]: statistics over rows < row_len[b].
template <typename T>
void grn_inplace(cudaStream_t st, T* h, int batches, int rows, int C, const int* row_len, const float* gamma,
                 const float* beta, float* scratch /* [batches, 2*C] */);

// TextEmbedding front (dit.py:86-115, unett.py:54-83): token lookup (+1 shift, filler 0), optional sinus position
// table, optional padding mask.  text int64 [B, nt]; branches: b2 < B cond, b2 >= B uncond (all-zero ids) when two.
void text_lookup(cudaStream_t st, const long long* text, int B, int nt, int branches_uncond_from, int batches,
                 int rows, int td, const int* row_len, const float* table, const float* inv_freq /* null: no pos */,
                 int mask_padding, float* out, unsigned char* keep, int table_rows, int* bad_id /* set to 1 on an id
                 outside the table (the reference's nn.Embedding raises IndexError); may be null */);
void mask_rows(cudaStream_t st, float* x, long long M, int d, const unsigned char* keep);

// A_in[m, 100:200) = cond (zero for the uncond branch), A_in[m, 200:200+td) = text embedding.
void build_a_in(cudaStream_t st, float* a_in, long long ld, int B, int batches, int rows, int mel, int td,
                const float* cond, const float* text_emb, int uncond_from);
// v holds `v_lead_rows` extra leading rows per batch element (UNetT's time token) that are skipped.
// y (+)= ...; also scatters y into the x columns of A_in for every branch and, optionally, into the trajectory.
void cfg_euler(cudaStream_t st, float* y, const float* v, int v_lead_rows, float* a_in, long long ld, int B, int branches, int rows,
               int mel, float dt, float cfg, float* traj_slot);

// out = masked cast of x; copy (optional) = x unchanged (seeds the residual stream for the conv pos-emb epilogue)
// Fused CFG + Euler update of y and x-half of the input projection: hproj[br] = hconst[br] + y W_x^T (null hproj: update
// only).  The new state goes to y_next (column tiles of a row tile run in different CTAs).  v == null leaves y unchanged (first evaluation / f5b_forward).
// Flow-matching training target (cfm.py:271-282): phi = (1 - t_b) x0 + t_b x1, cond = span ? 0 : x1.
void cfm_mix(cudaStream_t st, const float* x1, const float* x0, const float* t_dev, const unsigned char* span, int B,
             int rows, int mel, float* phi, float* cond);
// mean over the span frames of (pred - (x1 - x0))^2 (cfm.py:298-300); fixed reduction order.  partials: [2 * 256] doubles.
void masked_mse(cudaStream_t st, const float* pred, const float* x1, const float* x0, const unsigned char* span,
                long long rows_total, int mel, double* partials, float* loss);
void euler_xproj(cudaStream_t st, const float* y, float* y_next, const float* v, int v_lead_rows, int B, int branches, int rows, int mel,
                 float dt, float cfg, float* traj_slot, const float* Wx, long long ldw, const float* hconst,
                 float* hproj, int d);

template <typename T>
void cast_mask(cudaStream_t st, const float* x, T* out, int batches, int rows, int d, const int* row_len,
               float* copy = nullptr);

void rope_table(cudaStream_t st, float* cs, int rows, const float* inv_freq32);
void sinus_time(cudaStream_t st, const float* t, int n, const float* freq128, float* out /* [n,256] */);
void f32_to_bf16(cudaStream_t st, const float* in, bf16* out, long long n);
void fill_zero(cudaStream_t st, void* p, size_t bytes);

// UNetT helpers (unett.py:245, 261-270)
void unett_prepend(cudaStream_t st, const float* temb, const float* h, float* x, int batches, int rows, int d,
                   int temb_bmod);
template <typename T>
void concat2(cudaStream_t st, const float* a, const float* b, T* out, long long M, int d);

}  // namespace f5b
