// Backward kernels of the flow-matching training step, stage 1 (declarations; see train.cu).
#pragma once
#include "common.cuh"

namespace f5b {

// C[i,j] (+)= sum_k A[i*sa0 + k*sa1] * B[k*sb0 + j*sb1]
void sgemm_strided(cudaStream_t st, const float* A, long long sa0, long long sa1, const float* B, long long sb0,
                   long long sb1, float* C, long long ldc, int M, int N, int K, bool accumulate);
// scratch for the split-K form of skinny products (M <= 64, K >= 1024): >= 64 * M * N floats to use all 64 slices
void sgemm_set_scratch(float* p, size_t floats);
float* colsum_scratch(size_t* floats);  // the buffer handed to colsum_set_scratch (shared by the depthwise-conv wgrad)
// out[b*out_bstride + j] = sum_{r in batch b} X[r,j] * (Y ? Y[r,j] : 1)
// dY [R, C] fp32 -> bf16 copy, bf16 transpose [C, ldT] and column sums in one pass (false: shape not supported)
bool prep_dy16(cudaStream_t st, const float* src, long long R, int C, void* dst16, void* dstT, long long ldT, float* part,
               size_t part_floats, float* colsum_out);
// out = base + v[b(m), :] * x   and   dx = resid + LayerNorm-backward(dy)   (fused residual updates)
void rowscale_add(cudaStream_t st, const float* base, const float* x, const float* v, long long vec_bstride,
                  int rows_per_batch, long long M, int cols, float* out);
void ln_mod_bwd_resid(cudaStream_t st, const float* dy, const float* xhat, const float* rstd, long long M, int d,
                      const float* scale, long long vec_bstride, int rows_per_batch, const float* resid, float* dx);
void colsum_batch_pair(cudaStream_t st, const float* X, const float* Y, int batches, int rows_per_batch, int cols,
                       float* out_xy, float* out_x, long long out_bstride);
void colsum_batch(cudaStream_t st, const float* X, const float* Y, int batches, int rows_per_batch, int cols, float* out,
                  long long out_bstride);
// scratch for the two-level form (>= 128 * cols floats; per calling thread); null switches it off
void colsum_set_scratch(float* p, size_t floats);
void ln_mod_fwd_save(cudaStream_t st, const float* x, long long M, int d, const float* scale, const float* shift,
                     long long vec_bstride, int rows_per_batch, float eps, float* y, float* xhat, float* rstd);
void ln_mod_bwd(cudaStream_t st, const float* dy, const float* xhat, const float* rstd, long long M, int d,
                const float* scale, long long vec_bstride, int rows_per_batch, float* dx, bool accumulate);
void gelu_tanh_fwd(cudaStream_t st, const float* u, float* h, long long n);
void gelu_tanh_bwd(cudaStream_t st, const float* u, const float* dh, float* du, long long n);
void silu_bwd(cudaStream_t st, const float* z, const float* dy, float* dz, long long n);
void rowscale(cudaStream_t st, const float* x, const float* v, long long vec_bstride, int rows_per_batch,
              const int* row_len, long long M, int cols, float* out);
void axpy(cudaStream_t st, float* y, const float* x, float a, long long n);
void add_bias(cudaStream_t st, float* y, const float* bias, long long M, int cols);
void rope_qk(cudaStream_t st, float* qkv, long long M, int inner, int rows_per_batch, int rope_heads, const float* cs,
             float dir);
void mse_grad(cudaStream_t st, const float* pred, const float* x1, const float* x0, const unsigned char* span,
              long long rows, int mel, float* cnt_scratch, float* dpred);
void attention_backward(cudaStream_t st, const float* qkv, const float* o, const float* d_o, int batches, int rows,
                        int heads, const int* kv_len, float scale, float* lse, float* D, float* dqkv);

}  // namespace f5b

namespace f5b {
// embedding stage (train_embed.cu)
void mish_fwd_mask(cudaStream_t st, const float* z, float* y, int batches, int rows, int cols, const int* row_len);
void mish_bwd_mask(cudaStream_t st, const float* z, const float* dy, float* dz, int batches, int rows, int cols,
                   const int* row_len);
void gelu_erf_fwd(cudaStream_t st, const float* u, float* a, long long n);
void gelu_erf_bwd(cudaStream_t st, const float* u, const float* da, float* du, long long n);
void gconv_fwd(cudaStream_t st, const float* x, const float* w, const float* bias, int batches, int rows, int C, int groups,
               int taps, const int* row_len, float* z);
void gconv_dgrad(cudaStream_t st, const float* dz, const float* w, int batches, int rows, int C, int groups, int taps,
                 const int* row_len, float* dx);
void gconv_wgrad(cudaStream_t st, const float* dz, const float* x, int batches, int rows, int C, int groups, int taps,
                 const int* row_len, float* dw);
void grn_fwd(cudaStream_t st, const float* a, int batches, int rows, int C, const int* row_len, const float* gamma,
             const float* beta, float* stats, float* mu, float* g);
void grn_bwd(cudaStream_t st, const float* dg, const float* a, int batches, int rows, int C, const int* row_len,
             const float* gamma, float* stats, const float* mu, float* part, float* da, float* dgamma, float* dbeta);
void embed_bwd(cudaStream_t st, const float* dte, const long long* text, int B, int nt, int drop_text, int rows, int td,
               const int* row_len, int mask_padding, int masked_fill_applies, int table_rows, float* dtable);
void conv_w_flipT_bf16(cudaStream_t st, const float* w, void* out, int C, int groups, int taps);
void vec_add_const(cudaStream_t st, const float* v, float k, int n, float* out);
void col_slice(cudaStream_t st, const float* x, long long ldx, int off, int w, long long rows, float* y, long long ldy,
               int yoff);
}  // namespace f5b

namespace f5b {
// 16-bit operands for the tensor-core tier of the backward GEMMs (train.cu)
void cast16(cudaStream_t st, const float* src, long long lds, long long R, int C, void* dst, long long ldd, int fp16);
void transpose_cast16(cudaStream_t st, const float* src, long long lds, long long R, int C, void* dst, long long ldd,
                      int fp16);
void widen_bf16(cudaStream_t st, const void* src, long long n, float* dst);
}  // namespace f5b
