// Backward kernels of the flow-matching training step, stage 1 (declarations; see train.cu).
#pragma once
#include "common.cuh"

namespace f5b {

// C[i,j] (+)= sum_k A[i*sa0 + k*sa1] * B[k*sb0 + j*sb1]
void sgemm_strided(cudaStream_t st, const float* A, long long sa0, long long sa1, const float* B, long long sb0,
                   long long sb1, float* C, long long ldc, int M, int N, int K, bool accumulate);
// out[b*out_bstride + j] = sum_{r in batch b} X[r,j] * (Y ? Y[r,j] : 1)
void colsum_batch(cudaStream_t st, const float* X, const float* Y, int batches, int rows_per_batch, int cols, float* out,
                  long long out_bstride);
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
