// Optimiser step of the training loop (reference model/trainer.py:118-131 AdamW, :387 clip_grad_norm_, :395 EMA update)
// over FLAT fp32 buffers: one fused pass for the update, a fixed-tree double-precision reduction for the gradient norm.
// torch.optim.AdamW semantics (decoupled weight decay, bias correction).
#include "plan.h"

#include <mutex>
#include <string>

namespace f5b {

DeviceCtx open_device(int device, int mode);

// `gscale` folds the data-parallel average and the clip coefficient into the gradient read
__global__ void __launch_bounds__(256)
adamw_kernel(float* __restrict__ p, const float* __restrict__ g, float* __restrict__ m, float* __restrict__ v,
             float* __restrict__ ema, long long n, float lr, float b1, float b2, float eps, float wd, float bc1,
             float bc2_sqrt, float gscale, float ema_decay) {
  const long long i = (long long)blockIdx.x * 256 + threadIdx.x;
  if (i >= n) return;
  const float gi = g[i] * gscale;
  const float mi = b1 * m[i] + (1.0f - b1) * gi;
  const float vi = b2 * v[i] + (1.0f - b2) * gi * gi;
  m[i] = mi;
  v[i] = vi;
  float pi = p[i] * (1.0f - lr * wd);
  pi -= (lr / bc1) * mi / (sqrtf(vi) / bc2_sqrt + eps);
  p[i] = pi;
  if (ema != nullptr) ema[i] = ema_decay * ema[i] + (1.0f - ema_decay) * pi;
}

static constexpr int SQ_BLOCKS = 1024;
__global__ void __launch_bounds__(256)
sumsq_partial_kernel(const float* __restrict__ x, long long n, double* __restrict__ part) {
  __shared__ double red[256];
  double acc = 0.0;
  for (long long i = (long long)blockIdx.x * 256 + threadIdx.x; i < n; i += (long long)SQ_BLOCKS * 256)
    acc += (double)x[i] * (double)x[i];
  red[threadIdx.x] = acc;
  __syncthreads();
  for (int o = 128; o > 0; o >>= 1) {
    if ((int)threadIdx.x < o) red[threadIdx.x] += red[threadIdx.x + o];
    __syncthreads();
  }
  if (threadIdx.x == 0) part[blockIdx.x] = red[0];
}
__global__ void sumsq_final_kernel(const double* __restrict__ part, double* __restrict__ out) {
  double acc = 0.0;
  for (int i = 0; i < SQ_BLOCKS; ++i) acc += part[i];
  *out = acc;
}

// ---- allocation-free variant (f5b_op_sumsq_static) -------------------------------------------------------------
// The 1024 block partials live in this file-scope device array instead of a caller-supplied scratch buffer, so the
// entry point needs no device allocation from anyone (callers under CUDA-graph capture, tight serving loops, harnesses
// that forbid cudaMalloc).  Every GPU that loads the library holds the 8 KB until exit.  In HOST code the name
// `g_sumsq_scratch` is only a stand-in for the symbol, not a device address: the host resolves it once per device with
// cudaGetSymbolAddress (device current) and passes that pointer to the kernels.  All calls share the array (it keeps
// the last call's partials): calls that overlap on the GPU -- two streams, two host threads -- race on it.
static constexpr long long SQ_STATIC_MAX_N = 1LL << 31;  // 2 Gi elements = 8 GiB of fp32
static __device__ __align__(256) double g_sumsq_scratch[SQ_BLOCKS];
static constexpr int SQ_MAX_DEVICES = 64;
static double* g_sumsq_scratch_ptr[SQ_MAX_DEVICES] = {};
static std::mutex g_sumsq_mu;

// Device address of the scratch array on `device` (resolved once per device, with that device current).
static double* sumsq_scratch_on(int device) {
  F5B_REQUIRE(device >= 0 && device < SQ_MAX_DEVICES, "device index beyond the scratch table");
  std::lock_guard<std::mutex> lock(g_sumsq_mu);
  if (g_sumsq_scratch_ptr[device] == nullptr) {
    int n = 0;
    F5B_CHECK_CUDA(cudaGetDeviceCount(&n));
    F5B_REQUIRE(device < n, "device index out of range");
    F5B_CHECK_CUDA(cudaSetDevice(device));
    double* ptr = nullptr;
    F5B_CHECK_CUDA(cudaGetSymbolAddress((void**)&ptr, g_sumsq_scratch));
    F5B_REQUIRE(ptr != nullptr, "cudaGetSymbolAddress returned a null scratch address");
    g_sumsq_scratch_ptr[device] = ptr;
  }
  return g_sumsq_scratch_ptr[device];
}

}  // namespace f5b

using namespace f5b;

extern "C" {

// f5b_op_sumsq without a scratch argument and without any allocation (see g_sumsq_scratch above).  x, out: device
// pointers; n <= 2^31.  Nothing in the scratch has to start at zero: every one of the 1024 blocks overwrites its
// partial on every call.  Asynchronous on `stream`; calls that overlap on the GPU race on the shared scratch.
int f5b_op_sumsq_static(const float* x, int64_t n, double* out, int32_t device, void* stream) {
  try {
    F5B_REQUIRE(x && out && n > 0, "bad argument");
    if (n > SQ_STATIC_MAX_N) {
      set_error("f5b_op_sumsq_static: n = " + std::to_string((long long)n) + " exceeds the maximum of " +
                std::to_string(SQ_STATIC_MAX_N) + " elements of the allocation-free variant (use f5b_op_sumsq)");
      return F5B_ERR_INVALID;
    }
    double* scratch = sumsq_scratch_on(device);
    F5B_CHECK_CUDA(cudaSetDevice(device));
    cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
    sumsq_partial_kernel<<<SQ_BLOCKS, 256, 0, st>>>(x, n, scratch);
    F5B_POST_LAUNCH();
    sumsq_final_kernel<<<1, 1, 0, st>>>(scratch, out);
    F5B_POST_LAUNCH();
  } catch (const f5b::Err& e) {
    return e.code;
  }
  return F5B_OK;
}

// sum_i x[i]^2 in double precision -> *out (device pointer, 1 double); scratch: device, >= 1024 doubles.  Asynchronous.
int f5b_op_sumsq(const float* x, int64_t n, double* scratch, double* out, int32_t device, void* stream) {
  try {
    F5B_REQUIRE(x && scratch && out && n > 0, "bad argument");
    open_device(device, F5B_MODE_FP32);
    cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
    sumsq_partial_kernel<<<SQ_BLOCKS, 256, 0, st>>>(x, n, scratch);
    F5B_POST_LAUNCH();
    sumsq_final_kernel<<<1, 1, 0, st>>>(scratch, out);
    F5B_POST_LAUNCH();
  } catch (const f5b::Err& e) {
    return e.code;
  }
  return F5B_OK;
}

// One AdamW step over flat device buffers (param, grad, exp_avg, exp_avg_sq, optional EMA copy).  `step` counts from 1.
int f5b_op_adamw(float* param, const float* grad, float* exp_avg, float* exp_avg_sq, float* ema, int64_t n, float lr,
                 float beta1, float beta2, float eps, float weight_decay, int32_t step, float grad_scale,
                 float ema_decay, int32_t device, void* stream) {
  try {
    F5B_REQUIRE(param && grad && exp_avg && exp_avg_sq && n > 0 && step >= 1, "bad argument");
    open_device(device, F5B_MODE_FP32);
    cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
    const float bc1 = 1.0f - powf(beta1, (float)step), bc2 = 1.0f - powf(beta2, (float)step);
    adamw_kernel<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(param, grad, exp_avg, exp_avg_sq, ema, n, lr, beta1, beta2,
                                                            eps, weight_decay, bc1, sqrtf(bc2), grad_scale, ema_decay);
    F5B_POST_LAUNCH();
  } catch (const f5b::Err& e) {
    return e.code;
  }
  return F5B_OK;
}

}  // extern "C"
