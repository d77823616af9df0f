#pragma once
#include <cuda.h>
#include <cuda_runtime.h>
#include "common.cuh"

namespace f5b {

struct TcGemmShape {
  int rows_per_batch;
  int batches;
  int tiles_m_per_batch;
  int tiles_n;
  int N;
  int num_kb;
  int conv;
  int conv_kb_per_tap;
  int conv_pad;
  int conv_cg;
  int splits;      // split-K factor (residual epilogue only: partial sums meet in the L2 reduce-add)
  unsigned idesc;  // tcgen05 instruction descriptor (operand formats differ between the bf16 and fp16 variants)
};

// Grouped 1-D convolution expressed as a GEMM: K index = tap * cg + ci, A column = group * cg + ci,
// A row = output row + tap - pad (zero outside [0, rows_per_batch)).
struct TcConv {
  int taps;      // kernel size (31)
  int pad;       // 15
  int cg;        // channels per group
  int channels;  // total input channels (width of A)
  int fp16;      // operands are IEEE fp16 instead of bf16 (same tensor-core rate, 3 more mantissa bits)
};

void make_tmap(CUtensorMap* out, const void* base, int rank, const uint64_t* dims, const uint64_t* strides_bytes,
               const uint32_t* box, int elem_type /* OUT_F32 / OUT_BF16 / OUT_F16 */);
void make_tmap_bf16(CUtensorMap* out, const void* base, int rank, const uint64_t* dims, const uint64_t* strides_bytes,
                    const uint32_t* box, int fp16 = 0);

// bf16 tensor-core GEMM: out = epilogue(A[rows, K] * W[N, K]^T).  A rows are [batches][rows_per_batch] with a
// common leading dimension lda (elements).
void tc_gemm(cudaStream_t st, int sm_count, const bf16* A, long long lda, int rows_per_batch, int batches,
             const bf16* W, long long ldw, int N, int K, const EpiParams& ep, const TcConv* conv, int force_bn,
             int force_splits = 0, int fp16 = 0 /* operands are IEEE fp16 */);

// fp32 CUDA-core GEMM with the same epilogue (the correctness mode; also used for the tiny per-sample tables).
void simt_gemm(cudaStream_t st, const float* A, long long lda, long long M, const float* W, long long ldw, int N,
               int K, const EpiParams& ep);

// fp32 grouped conv (k taps, groups) + epilogue; x [batches, rows, channels], w rearranged [groups][taps][cg_in][cg_out].
void simt_group_conv(cudaStream_t st, const float* x, int batches, int rows, int channels, const float* w_r, int taps,
                     int pad, int groups, const EpiParams& ep);

// bf16 tensor-core attention (tcgen05) and fp32 CUDA-core attention.  qkv [batches*rows, 3*heads*64]; out
// [batches*rows, heads*64].  kv_len[b] (or rows when null) keys take part.
void tc_attention(cudaStream_t st, const bf16* qkv, bf16* out, int batches, int rows, int heads, const int* kv_len,
                  float scale, int out_f16 = 0 /* output written as IEEE fp16 (operand of the next GEMM) */,
                  float* lse2_out = nullptr /* [batches, heads, rows] log2-domain log-sum-exp (training) */);
// Backward of the same attention on the tensor cores (tc_attn_bwd.cu).  qkv bf16 as above, d_o bf16 [batches*rows,
// heads*64], lse2 from the forward, D[b,h,r] = sum_d dO O; dqkv fp32 [batches*rows, 3*heads*64] receives dq | dk | dv.
void tc_attention_backward(cudaStream_t st, const bf16* qkv, const bf16* d_o, const float* lse2, const float* D,
                           int batches, int rows, int heads, const int* kv_len, float scale, float* dqkv);
// d_o16 = bf16(d_o), D[b,h,r] = sum_d d_o[r, h*64+d] * o[r, h*64+d]   (one pass over both fp32 inputs)
void attn_bwd_prep(cudaStream_t st, const float* o, const float* d_o, int batches, int rows, int heads, bf16* d_o16,
                   float* D);
void tc_attn_bwd_init_kernels();
void simt_attention(cudaStream_t st, const float* qkv, float* out, int batches, int rows, int heads, const int* kv_len,
                    float scale);

void ktrace_set_tc_gemm(unsigned long long* p);  // per-translation-unit trace pointers (no-ops without F5B_KTRACE)
void ktrace_set_tc_attn(unsigned long long* p);
void ktrace_set_rowwise(unsigned long long* p);
void attn_trace_enable(int on);
void attn_trace_read(long long* host /* [3][64][8] clock64 stamps */);
void tc_init_kernels();       // attention
void tc_gemm_init_kernels();  // GEMM
void simt_init_kernels();

}  // namespace f5b
