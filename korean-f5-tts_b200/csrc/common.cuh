// Shared device/host helpers for libf5b200 (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <cuda_bf16.h>
#include <cuda_fp16.h>
#include <cstdint>
#include <cstdio>
#include <string>

#include "../../include/f5b200.h"

namespace f5b {

typedef __nv_bfloat16 bf16;
// 16-bit type of the linear-layer operands (activations and weights) in the tensor-core mode: IEEE fp16.
// tcgen05 kind::f16 contracts fp16 and bf16 at the same rate; fp16's 3 extra mantissa bits are what keeps the mode
// inside the 2e-2 mel tolerance at full depth; the attention kernel's q/k/v/P stay bf16.
typedef __half act16;

// ---------------------------------------------------------------------------------------------
// error handling: every C-ABI entry returns 0 or a negative code and leaves a thread-local message
// ---------------------------------------------------------------------------------------------
// error codes: F5B_OK / F5B_ERR_* from include/f5b200.h

void set_error(const std::string& msg);
const char* get_error();

struct Err {
  int code;
};

void count_launch(long long n = 1);  // kernel-launch accounting (f5b_launch_count)
long long launch_count();
void set_last_call(long long n);
long long last_call();

#define F5B_CHECK_CUDA(expr)                                                                      \
  do {                                                                                            \
    cudaError_t _e = (expr);                                                                      \
    if (_e != cudaSuccess) {                                                                      \
      f5b::set_error(std::string(#expr) + " -> " + cudaGetErrorString(_e) + " at " + __FILE__ +   \
                     ":" + std::to_string(__LINE__));                                             \
      throw f5b::Err{F5B_ERR_CUDA};                                                          \
    }                                                                                             \
  } while (0)

bool sync_debug();  // F5B_SYNC_DEBUG=1: synchronise after every launch so a device fault names its kernel

#define F5B_POST_LAUNCH()                                                                        \
  do {                                                                                            \
    F5B_CHECK_CUDA(cudaGetLastError());                                                           \
    if (f5b::sync_debug()) F5B_CHECK_CUDA(cudaDeviceSynchronize());                               \
    f5b::count_launch(1);                                                                         \
  } while (0)

#define F5B_REQUIRE(cond, msg)                                                                    \
  do {                                                                                            \
    if (!(cond)) {                                                                                \
      f5b::set_error(std::string(msg) + " [" #cond "] at " + __FILE__ + ":" +                     \
                     std::to_string(__LINE__));                                                   \
      throw f5b::Err{F5B_ERR_INVALID};                                                       \
    }                                                                                             \
  } while (0)

// ---------------------------------------------------------------------------------------------
// epilogue shared by the SIMT (fp32) and tcgen05 (bf16) GEMMs
// ---------------------------------------------------------------------------------------------
enum : int { EPI_STORE = 0, EPI_RESID = 1 };
enum : int { OUT_F32 = 0, OUT_BF16 = 1, OUT_F16 = 2 };
enum : int { ACT_NONE = 0, ACT_GELU_TANH = 1, ACT_GELU_ERF = 2, ACT_MISH = 3, ACT_SILU = 4 };

struct EpiParams {
  int kind = EPI_STORE;
  int act = ACT_NONE;
  int out_type = 0;               // STORE target element type: OUT_F32 / OUT_BF16 / OUT_F16
  int n_valid = 0;                // columns >= n_valid are dropped
  const float* bias = nullptr;    // [N] or null
  void* out = nullptr;            // STORE: out[m*ldo + n] = act(v) (+ addend)
  long long ldo = 0;
  const float* addend = nullptr;  // STORE: fp32 [M, ld_add] added after the activation
  long long ld_add = 0;
  float* resid = nullptr;         // RESID: resid[m*ldr + n] += g * v
  long long ldr = 0;
  const float* gate = nullptr;    // g = gate[(b % gate_bmod) * gate_bstride + n], null -> 1
  long long gate_bstride = 0;
  int gate_bmod = 1;
  int rows_per_batch = 1 << 30;   // b = m / rows_per_batch, pos = m % rows_per_batch
  const int* row_len = nullptr;   // [batches]: rows with pos >= row_len[b] store 0 (STORE) / are skipped (RESID)
  int row_skip = 0;               // rows with pos < row_skip are never written (UNetT time token on proj_out)
  int rope_cols = 0;              // columns < rope_cols hold q|k heads; head = (col % rope_period)/64
  int rope_period = 1;
  int rope_heads = 0;             // heads < rope_heads are rotated
  const float* rope_cs = nullptr; // [rows_per_batch, 64] interleaved (cos_j, sin_j), j < 32
};

bool pdl_enabled(int kernel_id = 0);  // programmatic dependent launch for the per-step kernel chain (F5B_NO_PDL=1 switches it off)

#if defined(__CUDACC__)

// Programmatic dependent launch.  Every kernel of the per-step chain calls pdl_trigger() first (its successor may
// be scheduled as soon as all of this grid's CTAs are resident) and pdl_wait() before it touches anything an earlier
// kernel of the chain writes (or writes anything an earlier kernel reads); what runs before pdl_wait() -- barrier
// init, TMEM allocation, descriptor prefetch, loads of weights / per-call tables -- overlaps the predecessor's tail.
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void pdl_trigger() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }

// In-situ kernel timeline (debug builds only: `F5B_KTRACE=1 python build.py --force` adds -DF5B_KTRACE).  Thread 0 of
// every CTA of the per-step chain records %globaltimer at kernel entry, after the programmatic-dependency wait and at
// exit; f5b_debug_ktrace() returns the records, tools/ktrace.py turns them into a per-launch timeline (launch gap,
// ramp, tail) of the graph-replayed sampling loop -- the view nsys would give.  Without the flag the macros vanish.
#if defined(F5B_KTRACE)
static __device__ unsigned long long* g_ktrace;  // per translation unit; [0] record count, [1] capacity, records from [4]
__device__ __forceinline__ unsigned long long ktrace_now() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  return t;
}
#define KT_DECL __shared__ unsigned long long kt_stamp[6];
#define KT_MARK(i) do { kt_stamp[i] = ktrace_now(); } while (0)  /* kernel-specific phase stamps 2..5 */
#define KT_BEGIN()  do { if (threadIdx.x == 0) kt_stamp[0] = ktrace_now(); } while (0)
#define KT_WAITED() do { if (threadIdx.x == 0) kt_stamp[1] = ktrace_now(); } while (0)
#define KT_END(kid)                                                                                         \
  do {                                                                                                      \
    if (threadIdx.x == 0 && g_ktrace) {                                                                     \
      unsigned long long* kt_p = g_ktrace;                                                                  \
      unsigned long long kt_i = atomicAdd(kt_p, 1ULL);                                                      \
      if (kt_i < kt_p[1]) {                                                                                 \
        unsigned long long* r = kt_p + 4 + 8 * kt_i;                                                        \
        r[0] = kt_stamp[0]; r[1] = kt_stamp[1]; r[2] = ktrace_now();                                        \
        r[3] = ((unsigned long long)(kid) << 32) | (blockIdx.x + gridDim.x * (blockIdx.y + gridDim.y * blockIdx.z)); \
        r[4] = kt_stamp[2]; r[5] = kt_stamp[3]; r[6] = kt_stamp[4]; r[7] = kt_stamp[5];                    \
      }                                                                                                     \
    }                                                                                                       \
  } while (0)
#define KT_DEFINE_SETTER(name) \
  void name(unsigned long long* p) { F5B_CHECK_CUDA(cudaMemcpyToSymbol(g_ktrace, &p, sizeof(p))); }
#else
#define KT_DECL
#define KT_BEGIN() do { } while (0)
#define KT_WAITED() do { } while (0)
#define KT_MARK(i) do { } while (0)
#define KT_END(kid) do { } while (0)
#define KT_DEFINE_SETTER(name) \
  void name(unsigned long long*) {}
#endif

template <typename... P, typename... A>
inline void launch_pdl_cluster(int kernel_id, int cluster_x, void (*kern)(P...), dim3 grid, dim3 block, size_t smem,
                               cudaStream_t st, A&&... args) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = grid;
  cfg.blockDim = block;
  cfg.dynamicSmemBytes = smem;
  cfg.stream = st;
  cudaLaunchAttribute at[2];
  at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  at[0].val.programmaticStreamSerializationAllowed = pdl_enabled(kernel_id) ? 1 : 0;
  cfg.numAttrs = 1;
  if (cluster_x > 1) {  // thread-block cluster along x (CTA pairs of the tcgen05 cta_group::2 kernels)
    at[1].id = cudaLaunchAttributeClusterDimension;
    at[1].val.clusterDim.x = (unsigned)cluster_x;
    at[1].val.clusterDim.y = 1;
    at[1].val.clusterDim.z = 1;
    cfg.numAttrs = 2;
  }
  cfg.attrs = at;
  F5B_CHECK_CUDA(cudaLaunchKernelEx(&cfg, kern, P(args)...));
}
template <typename... P, typename... A>
inline void launch_pdl_id(int kernel_id, void (*kern)(P...), dim3 grid, dim3 block, size_t smem, cudaStream_t st,
                          A&&... args) {
  launch_pdl_cluster(kernel_id, 1, kern, grid, block, smem, st, static_cast<A&&>(args)...);
}

__device__ __forceinline__ float act_apply(int act, float x) {
  switch (act) {
    case ACT_GELU_TANH: {
      const float k = 0.7978845608028654f;
      float u = k * (x + 0.044715f * x * x * x);
      return 0.5f * x * (1.0f + tanhf(u));
    }
    case ACT_GELU_ERF:
      return 0.5f * x * (1.0f + erff(x * 0.7071067811865476f));
    case ACT_MISH: {
      float sp = x > 20.0f ? x : log1pf(expf(x));
      return x * tanhf(sp);
    }
    case ACT_SILU:
      return x / (1.0f + expf(-x));
    default:
      return x;
  }
}

__device__ __forceinline__ uint32_t pack_bf16x2(float a, float b) {
  __nv_bfloat162 h = __floats2bfloat162_rn(a, b);
  return *reinterpret_cast<uint32_t*>(&h);
}

__device__ __forceinline__ uint32_t pack_f16x2(float a, float b) {
  __half2 h = __floats2half2_rn(a, b);
  return *reinterpret_cast<uint32_t*>(&h);
}

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}

__device__ __forceinline__ float ld_as_float(const float* p) { return *p; }
__device__ __forceinline__ float ld_as_float(const bf16* p) { return __bfloat162float(*p); }
__device__ __forceinline__ float ld_as_float(const __half* p) { return __half2float(*p); }
__device__ __forceinline__ void st_from_float(float* p, float v) { *p = v; }
__device__ __forceinline__ void st_from_float(bf16* p, float v) { *p = __float2bfloat16_rn(v); }
__device__ __forceinline__ void st_from_float(__half* p, float v) { *p = __float2half_rn(v); }

// Apply the epilogue to NC (multiple of 4) consecutive accumulator columns [n0, n0+NC) of global row m.
// n0 must be a multiple of 4 (RoPE pairs and the vector stores rely on it).
template <int NC>
__device__ __forceinline__ void epi_apply(const EpiParams& p, long long m, int n0, float* v) {
  const int b = (int)(m / p.rows_per_batch);
  const int pos = (int)(m - (long long)b * p.rows_per_batch);
  if (pos < p.row_skip) return;
  const long long mo = m - (long long)(b + 1) * p.row_skip;  // output row when leading rows are dropped
  const bool row_dead = (p.row_len != nullptr) && (pos >= p.row_len[b]);
  if (p.kind == EPI_RESID && row_dead) return;
  if (n0 >= p.n_valid) return;

  if (p.bias != nullptr) {
#pragma unroll
    for (int j = 0; j < NC; j += 4) {
      if (n0 + j + 4 <= p.n_valid) {
        float4 bv = __ldg(reinterpret_cast<const float4*>(p.bias + n0 + j));
        v[j] += bv.x; v[j + 1] += bv.y; v[j + 2] += bv.z; v[j + 3] += bv.w;
      } else {
#pragma unroll
        for (int q = 0; q < 4; ++q)
          if (n0 + j + q < p.n_valid) v[j + q] += __ldg(p.bias + n0 + j + q);
      }
    }
  }
  if (p.rope_cs != nullptr && n0 < p.rope_cols) {
    const float* cs = p.rope_cs + (long long)pos * 64;
#pragma unroll
    for (int j = 0; j < NC; j += 2) {
      int col = n0 + j;
      if (col < p.rope_cols) {
        int within = col % p.rope_period;
        if ((within >> 6) < p.rope_heads) {
          float2 t = __ldg(reinterpret_cast<const float2*>(cs + (within & 63)));
          float a = v[j], c = v[j + 1];
          v[j] = a * t.x - c * t.y;
          v[j + 1] = c * t.x + a * t.y;
        }
      }
    }
  }
  if (p.act != ACT_NONE) {
#pragma unroll
    for (int j = 0; j < NC; ++j) v[j] = act_apply(p.act, v[j]);
  }

  if (p.kind == EPI_RESID) {
    float* r = p.resid + m * p.ldr + n0;
    const float* g = p.gate ? p.gate + (long long)(b % p.gate_bmod) * p.gate_bstride + n0 : nullptr;
#pragma unroll
    for (int j = 0; j < NC; j += 4) {
      if (n0 + j + 4 <= p.n_valid) {
        float4 x = *reinterpret_cast<float4*>(r + j);
        if (g) {
          float4 gv = __ldg(reinterpret_cast<const float4*>(g + j));
          x.x += gv.x * v[j]; x.y += gv.y * v[j + 1]; x.z += gv.z * v[j + 2]; x.w += gv.w * v[j + 3];
        } else {
          x.x += v[j]; x.y += v[j + 1]; x.z += v[j + 2]; x.w += v[j + 3];
        }
        *reinterpret_cast<float4*>(r + j) = x;
      } else {
#pragma unroll
        for (int q = 0; q < 4; ++q)
          if (n0 + j + q < p.n_valid) r[j + q] += (g ? __ldg(g + j + q) : 1.0f) * v[j + q];
      }
    }
    return;
  }

  // EPI_STORE
  if (row_dead) {
#pragma unroll
    for (int j = 0; j < NC; ++j) v[j] = 0.0f;
  }
  if (p.addend != nullptr) {  // added even on dead rows: x = conv_pos(h) + h keeps h there (conv output is 0)
    const float* a = p.addend + mo * p.ld_add + n0;
#pragma unroll
    for (int j = 0; j < NC; j += 4) {
      if (n0 + j + 4 <= p.n_valid) {
        float4 av = *reinterpret_cast<const float4*>(a + j);
        v[j] += av.x; v[j + 1] += av.y; v[j + 2] += av.z; v[j + 3] += av.w;
      } else {
#pragma unroll
        for (int q = 0; q < 4; ++q)
          if (n0 + j + q < p.n_valid) v[j + q] += a[j + q];
      }
    }
  }
  if (p.out_type == OUT_BF16) {
    bf16* o = reinterpret_cast<bf16*>(p.out) + mo * p.ldo + n0;
#pragma unroll
    for (int j = 0; j < NC; j += 4) {
      if (n0 + j + 4 <= p.n_valid) {
        uint2 pk;
        pk.x = pack_bf16x2(v[j], v[j + 1]);
        pk.y = pack_bf16x2(v[j + 2], v[j + 3]);
        *reinterpret_cast<uint2*>(o + j) = pk;
      } else {
#pragma unroll
        for (int q = 0; q < 4; ++q)
          if (n0 + j + q < p.n_valid) o[j + q] = __float2bfloat16_rn(v[j + q]);
      }
    }
  } else if (p.out_type == OUT_F16) {
    __half* o = reinterpret_cast<__half*>(p.out) + mo * p.ldo + n0;
#pragma unroll
    for (int j = 0; j < NC; j += 4) {
      if (n0 + j + 4 <= p.n_valid) {
        uint2 pk;
        pk.x = pack_f16x2(v[j], v[j + 1]);
        pk.y = pack_f16x2(v[j + 2], v[j + 3]);
        *reinterpret_cast<uint2*>(o + j) = pk;
      } else {
#pragma unroll
        for (int q = 0; q < 4; ++q)
          if (n0 + j + q < p.n_valid) o[j + q] = __float2half_rn(v[j + q]);
      }
    }
  } else {
    float* o = reinterpret_cast<float*>(p.out) + mo * p.ldo + n0;
#pragma unroll
    for (int j = 0; j < NC; j += 4) {
      if (n0 + j + 4 <= p.n_valid) {
        *reinterpret_cast<float4*>(o + j) = make_float4(v[j], v[j + 1], v[j + 2], v[j + 3]);
      } else {
#pragma unroll
        for (int q = 0; q < 4; ++q)
          if (n0 + j + q < p.n_valid) o[j + q] = v[j + q];
      }
    }
  }
}

#endif  // __CUDACC__

}  // namespace f5b
