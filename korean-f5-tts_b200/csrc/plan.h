// Host-side runtime of libf5b200: weight store, workspace arena, backbone execution, CUDA-graph cache.
#pragma once
#include <map>
#include <memory>
#include <string>
#include <vector>

#include "../../include/f5b200.h"
#include "common.cuh"
#include "rowwise.h"
#include "tc_gemm.h"

namespace f5b {

struct DeviceCtx {
  int device = 0;
  int sm_count = 148;
  int mode = F5B_MODE_FP32;
};

// Owns cudaMalloc'd blocks; everything is released together.
// Replay: code that allocates a deterministic sequence of buffers (the plan's weight packing, the training step's
// workspace) can be run again over the SAME blocks -- begin_replay() rewinds, alloc() then hands out the existing
// blocks in order (a block that is too small is replaced by a larger one, further requests are appended), so a repeated
// call costs no cudaMalloc / cudaFree (each of which synchronises the device) and batches of changing shape settle on
// the largest sizes seen.
class Arena {
 public:
  ~Arena() { release(); }
  void* alloc(size_t bytes);
  template <typename T>
  T* alloc_n(size_t n) { return reinterpret_cast<T*>(alloc(n * sizeof(T))); }
  void release();
  size_t total() const { return total_; }
  void begin_replay() { replay_ = 0; }
  void end_replay() { replay_ = (size_t)-1; }
  size_t blocks() const { return blocks_.size(); }

 private:
  std::vector<void*> blocks_;
  std::vector<size_t> sizes_;
  size_t total_ = 0;
  size_t replay_ = (size_t)-1;  // index of the next block to hand out again; -1: not replaying
};

// A weight matrix usable by either contraction path: fp32 [N, K] always; bf16 [N, ldw] in bf16 mode.
struct WMat {
  const float* f32 = nullptr;
  const bf16* b16 = nullptr;  // 16-bit copy: IEEE fp16 bits when `fp16`, bf16 otherwise
  int fp16 = 0;
  int N = 0, K = 0;
  long long ld32 = 0, ld16 = 0;
};

class WeightStore {
 public:
  WeightStore(Arena* arena, const f5b_tensor* w, int n, cudaStream_t st);
  const float* get(const std::string& name, long long numel) const;  // throws F5B_ERR_MISSING
  bool has(const std::string& name) const { return map_.count(name) != 0; }

 private:
  struct Entry { float* p; long long numel; };
  std::map<std::string, Entry> map_;
};

// out = epilogue(A[M,K] W^T): dispatch on the plan's precision mode.  A is `float` (fp32 mode) or `bf16`.
void run_gemm(const DeviceCtx& ctx, cudaStream_t st, const void* A, long long lda, long long M, const WMat& w,
              const EpiParams& ep, int force_bn = 0);
WMat make_wmat(const DeviceCtx& ctx, Arena* arena, cudaStream_t st, const float* w32, int N, int K, int fp16 = 1);
WMat make_wmat_split3(const DeviceCtx& ctx, Arena* arena, cudaStream_t st, const float* w32, int N, int K);
// cat rows of several [Ni, K] fp32 matrices into one device matrix
float* concat_rows(Arena* arena, cudaStream_t st, const std::vector<std::pair<const float*, long long>>& parts);

void copy_in(void* dst, const void* src, size_t bytes, int mem, cudaStream_t st);
void copy_out(void* dst, const void* src, size_t bytes, int mem, cudaStream_t st);

}  // namespace f5b
