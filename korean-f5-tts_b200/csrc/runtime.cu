// Runtime plumbing: errors, arena, weight store, contraction dispatch, kernel-level C entry points.
#include <atomic>
#include <cstdlib>
#include <cmath>
#include <cstring>
#include <mutex>
#include <string>

#include "plan.h"
#include "train.h"

namespace f5b {

static thread_local std::string g_err;
void set_error(const std::string& msg) { g_err = msg; }
const char* get_error() { return g_err.c_str(); }

bool pdl_enabled(int kernel_id) {  // F5B_PDL_MASK (debug): bit set = that kernel family launches with PDL
  static const bool on = std::getenv("F5B_NO_PDL") == nullptr;
  static const int mask = std::getenv("F5B_PDL_MASK") ? std::atoi(std::getenv("F5B_PDL_MASK")) : ~0;
  return on && (kernel_id == 0 || (mask & kernel_id) != 0);
}
bool sync_debug() {
  static const bool on = std::getenv("F5B_SYNC_DEBUG") != nullptr;
  return on;
}

static std::atomic<long long> g_launches{0};
void count_launch(long long n) { g_launches.fetch_add(n, std::memory_order_relaxed); }
long long launch_count() { return g_launches.load(std::memory_order_relaxed); }
static thread_local long long g_last_call = 0;
void set_last_call(long long n) { g_last_call = n; }
long long last_call() { return g_last_call; }

void* Arena::alloc(size_t bytes) {
  void* p = nullptr;
  size_t b = (bytes + 255) & ~size_t(255);
  if (b == 0) b = 256;
  if (replay_ != (size_t)-1 && replay_ < blocks_.size()) {
    const size_t i = replay_++;
    if (sizes_[i] >= b) return blocks_[i];  // grow-only: a smaller request (a shorter batch) reuses the larger block
    F5B_CHECK_CUDA(cudaFree(blocks_[i]));   // the sequence needs more here than ever before: replace this block
    blocks_[i] = nullptr;
    total_ -= sizes_[i];
    sizes_[i] = 0;
    F5B_CHECK_CUDA(cudaMalloc(&p, b));
    blocks_[i] = p;
    sizes_[i] = b;
    total_ += b;
    return p;
  }
  F5B_CHECK_CUDA(cudaMalloc(&p, b));
  blocks_.push_back(p);
  sizes_.push_back(b);
  total_ += b;
  if (replay_ != (size_t)-1) replay_ = blocks_.size();
  return p;
}
void Arena::release() {
  for (void* p : blocks_)
    if (p) cudaFree(p);
  blocks_.clear();
  sizes_.clear();
  total_ = 0;
  replay_ = (size_t)-1;
}

WeightStore::WeightStore(Arena* arena, const f5b_tensor* w, int n, cudaStream_t st) {
  for (int i = 0; i < n; ++i) {
    F5B_REQUIRE(w[i].name != nullptr && w[i].data != nullptr && w[i].numel > 0, "bad weight entry");
    float* d = arena->alloc_n<float>((size_t)w[i].numel);
    F5B_CHECK_CUDA(cudaMemcpyAsync(d, w[i].data, (size_t)w[i].numel * 4,
                                   w[i].on_device ? cudaMemcpyDeviceToDevice : cudaMemcpyHostToDevice, st));
    map_[w[i].name] = Entry{d, w[i].numel};
  }
  F5B_CHECK_CUDA(cudaStreamSynchronize(st));
}
const float* WeightStore::get(const std::string& name, long long numel) const {
  auto it = map_.find(name);
  if (it == map_.end()) {
    set_error("missing weight '" + name + "'");
    throw Err{F5B_ERR_MISSING};
  }
  if (numel >= 0 && it->second.numel != numel) {
    set_error("weight '" + name + "' has " + std::to_string(it->second.numel) + " elements, expected " +
              std::to_string(numel));
    throw Err{F5B_ERR_INVALID};
  }
  return it->second.p;
}

template <typename T>
__global__ void pad_convert_kernel(const float* __restrict__ in, T* __restrict__ out, int K, long long ld,
                                   long long total) {
  long long i = (long long)blockIdx.x * 256 + threadIdx.x;
  if (i >= total) return;
  long long r = i / ld;
  int c = (int)(i - r * ld);
  st_from_float(out + i, c < K ? in[r * K + c] : 0.0f);
}

// Pseudo-random 16-bit operands for the kernel timers (f5b_op_time): constant data toggles few bits, and tensor-core
// power -- on a power-capped part, the clock -- depends on the data.  Uniform in [-amp, amp), hash of the index.
__global__ void fill_random16_kernel(uint16_t* __restrict__ out, long long n, float amp, uint32_t seed, int fp16) {
  long long i = (long long)blockIdx.x * 256 + threadIdx.x;
  if (i >= n) return;
  uint32_t h = (uint32_t)i * 2654435761u ^ seed;
  h ^= h >> 16; h *= 0x7feb352du; h ^= h >> 15; h *= 0x846ca68bu; h ^= h >> 16;
  const float v = ((h >> 8) * (1.0f / 8388608.0f) - 1.0f) * amp;
  if (fp16) {
    __half x = __float2half_rn(v);
    out[i] = *reinterpret_cast<uint16_t*>(&x);
  } else {
    bf16 x = __float2bfloat16_rn(v);
    out[i] = *reinterpret_cast<uint16_t*>(&x);
  }
}
static void fill_random16(cudaStream_t st, void* p, long long n, float amp, uint32_t seed, int fp16 = 0) {
  fill_random16_kernel<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(reinterpret_cast<uint16_t*>(p), n, amp, seed, fp16);
  F5B_POST_LAUNCH();
}
__global__ void fill_random32_kernel(float* __restrict__ out, long long n, float amp, uint32_t seed) {
  long long i = (long long)blockIdx.x * 256 + threadIdx.x;
  if (i >= n) return;
  uint32_t h = (uint32_t)i * 2654435761u ^ seed;
  h ^= h >> 16; h *= 0x7feb352du; h ^= h >> 15; h *= 0x846ca68bu; h ^= h >> 16;
  out[i] = ((h >> 8) * (1.0f / 8388608.0f) - 1.0f) * amp;
}

__global__ void widen_kernel(const bf16* __restrict__ in, float* __restrict__ o, long long n) {
  long long i = (long long)blockIdx.x * 256 + threadIdx.x;
  if (i < n) o[i] = __bfloat162float(in[i]);
}

WMat make_wmat(const DeviceCtx& ctx, Arena* arena, cudaStream_t st, const float* w32, int N, int K, int fp16) {
  WMat m;
  m.f32 = w32;
  m.N = N;
  m.K = K;
  m.ld32 = K;
  if (ctx.mode == F5B_MODE_BF16) {
    long long ld = (K + 7) / 8 * 8;
    bf16* p = arena->alloc_n<bf16>((size_t)N * ld);
    long long total = (long long)N * ld;
    if (fp16)
      pad_convert_kernel<__half><<<(unsigned)((total + 255) / 256), 256, 0, st>>>(w32, reinterpret_cast<__half*>(p), K,
                                                                                 ld, total);
    else
      pad_convert_kernel<bf16><<<(unsigned)((total + 255) / 256), 256, 0, st>>>(w32, p, K, ld, total);
    m.fp16 = fp16;
    F5B_POST_LAUNCH();
    m.b16 = p;
    m.ld16 = ld;
  }
  return m;
}

__global__ void split3_weight_kernel(const float* __restrict__ in, bf16* __restrict__ out, int K, long long total) {
  long long i = (long long)blockIdx.x * 256 + threadIdx.x;
  if (i >= total) return;
  long long r = i / K;
  int c = (int)(i - r * K);
  float v = in[i];
  bf16 hi = __float2bfloat16_rn(v);
  bf16 lo = __float2bfloat16_rn(v - __bfloat162float(hi));
  bf16* o = out + r * 3 * K;
  o[c] = hi;
  o[K + c] = hi;
  o[2 * K + c] = lo;
}

// [N, K] fp32 -> bf16 [N, 3K] = [W_hi | W_hi | W_lo], the partner of the bf16x3 activation layout (rowwise.h).
WMat make_wmat_split3(const DeviceCtx& ctx, Arena* arena, cudaStream_t st, const float* w32, int N, int K) {
  WMat m;
  m.f32 = w32;
  m.N = N;
  m.K = K;
  m.ld32 = K;
  if (ctx.mode == F5B_MODE_BF16) {
    F5B_REQUIRE(K % 8 == 0, "split3 weights need K % 8 == 0");
    bf16* p = arena->alloc_n<bf16>((size_t)N * 3 * K);
    long long total = (long long)N * K;
    split3_weight_kernel<<<(unsigned)((total + 255) / 256), 256, 0, st>>>(w32, p, K, total);
    F5B_POST_LAUNCH();
    m.b16 = p;
    m.ld16 = 3LL * K;
    m.K = 3 * K;
  }
  return m;
}

float* concat_rows(Arena* arena, cudaStream_t st, const std::vector<std::pair<const float*, long long>>& parts) {
  long long total = 0;
  for (auto& p : parts) total += p.second;
  float* out = arena->alloc_n<float>((size_t)total);
  long long off = 0;
  for (auto& p : parts) {
    F5B_CHECK_CUDA(cudaMemcpyAsync(out + off, p.first, (size_t)p.second * 4, cudaMemcpyDeviceToDevice, st));
    off += p.second;
  }
  return out;
}

void run_gemm(const DeviceCtx& ctx, cudaStream_t st, const void* A, long long lda, long long M, const WMat& w,
              const EpiParams& ep, int force_bn) {
  if (ctx.mode == F5B_MODE_BF16) {
    F5B_REQUIRE(w.b16 != nullptr, "bf16 weight missing");
    F5B_REQUIRE(M < (1LL << 31), "GEMM rows overflow");
    int rpb = (int)M, nb = 1;
    if (ep.row_skip > 0) {  // dropped leading rows are expressed as a TMA coordinate shift per batch element
      F5B_REQUIRE(M % ep.rows_per_batch == 0, "row_skip needs whole batch elements");
      rpb = ep.rows_per_batch;
      nb = (int)(M / ep.rows_per_batch);
    }
    tc_gemm(st, ctx.sm_count, reinterpret_cast<const bf16*>(A), lda, rpb, nb, w.b16, w.ld16, w.N, w.K, ep, nullptr,
            force_bn, 0, w.fp16);
  } else {
    simt_gemm(st, reinterpret_cast<const float*>(A), lda, M, w.f32, w.ld32, w.N, w.K, ep);
  }
}

void copy_in(void* dst, const void* src, size_t bytes, int mem, cudaStream_t st) {
  F5B_CHECK_CUDA(cudaMemcpyAsync(dst, src, bytes, mem == F5B_MEM_HOST ? cudaMemcpyHostToDevice : cudaMemcpyDeviceToDevice, st));
}
void copy_out(void* dst, const void* src, size_t bytes, int mem, cudaStream_t st) {
  F5B_CHECK_CUDA(cudaMemcpyAsync(dst, src, bytes, mem == F5B_MEM_HOST ? cudaMemcpyDeviceToHost : cudaMemcpyDeviceToDevice, st));
}

// Device properties are queried ONCE per device (cudaGetDeviceProperties costs milliseconds; every f5b_op_* entry and
// every optimiser step used to pay it twice) and the kernel attributes are set once per device as well.
struct DeviceInfo {
  std::atomic<bool> checked{false}, kernels_ready{false};  // read on the lock-free fast path, written under g_dev_mu
  int major = 0, minor = 0, sm_count = 0;
  std::string name;
};
static std::mutex g_dev_mu;
static DeviceInfo g_dev[64];

static DeviceInfo& device_check_throw(int device) {
  if (device >= 0 && device < 64 && g_dev[device].checked.load(std::memory_order_acquire) &&
      g_dev[device].major == 10) {  // fast path: the fields are written once, before `checked` is released
    F5B_CHECK_CUDA(cudaSetDevice(device));
    return g_dev[device];
  }
  int n = 0;
  cudaError_t e = cudaGetDeviceCount(&n);
  if (e != cudaSuccess || n == 0) {
    cudaGetLastError();
    set_error(std::string("no CUDA device: libf5b200 has no CPU fallback (") + cudaGetErrorString(e) + ")");
    throw Err{F5B_ERR_NO_DEVICE};
  }
  if (device < 0 || device >= n || device >= 64) {
    set_error("device index out of range");
    throw Err{F5B_ERR_NO_DEVICE};
  }
  std::lock_guard<std::mutex> lock(g_dev_mu);
  DeviceInfo& d = g_dev[device];
  if (!d.checked) {
    cudaDeviceProp p;
    F5B_CHECK_CUDA(cudaGetDeviceProperties(&p, device));
    d.major = p.major;
    d.minor = p.minor;
    d.sm_count = p.multiProcessorCount;
    d.name = p.name;
    d.checked.store(true, std::memory_order_release);
  }
  if (d.major != 10) {
    set_error(std::string("device '") + d.name + "' is sm_" + std::to_string(d.major) + std::to_string(d.minor) +
              "; libf5b200 is built for sm_100a only");
    throw Err{F5B_ERR_NO_DEVICE};
  }
  F5B_CHECK_CUDA(cudaSetDevice(device));
  return d;
}

DeviceCtx open_device(int device, int mode) {
  DeviceInfo& d = device_check_throw(device);
  DeviceCtx c;
  c.device = device;
  c.mode = mode;
  c.sm_count = d.sm_count;
  if (!d.kernels_ready) {  // per DEVICE: the dynamic shared-memory limits are attributes of a function on a device
    std::lock_guard<std::mutex> lock(g_dev_mu);
    if (!d.kernels_ready) {
      tc_init_kernels();
      tc_attn_bwd_init_kernels();
      tc_gemm_init_kernels();
      simt_init_kernels();
      d.kernels_ready.store(true, std::memory_order_release);
    }
  }
  return c;
}

}  // namespace f5b

using namespace f5b;
namespace f5b { __global__ void widen_kernel(const bf16*, float*, long long); }

#define F5B_API_BEGIN try {
#define F5B_API_END                                  \
  }                                                  \
  catch (const f5b::Err& e) { return e.code; }       \
  catch (const std::exception& e) {                  \
    f5b::set_error(std::string("exception: ") + e.what()); \
    return F5B_ERR_CUDA;                             \
  }                                                  \
  return F5B_OK;

namespace f5b { DeviceCtx open_device(int device, int mode); }

extern "C" {

const char* f5b_last_error(void) { return get_error(); }
int f5b_version(void) { return 100; }
int64_t f5b_launch_count(void) { return launch_count(); }
int64_t f5b_last_call_launches(void) { return last_call(); }

int f5b_device_check(int device) {
  F5B_API_BEGIN
  open_device(device, F5B_MODE_FP32);
  F5B_API_END
}

int f5b_op_gemm(int32_t mode, const float* A, const float* W, const float* bias, int32_t M, int32_t N, int32_t K,
                int32_t act, float* out, int32_t force_bn, int32_t device, void* stream) {
  F5B_API_BEGIN
  const int fp16 = mode == 2;  // kernel-level switch: 1 = bf16 operands, 2 = IEEE fp16 operands
  if (mode == 2) mode = F5B_MODE_BF16;
  DeviceCtx ctx = open_device(device, mode);
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  Arena arena;
  WMat w = make_wmat(ctx, &arena, st, W, N, K, fp16);
  EpiParams ep;
  ep.act = act;
  ep.bias = bias;
  ep.out = out;
  ep.ldo = N;
  ep.n_valid = N;
  if (mode == F5B_MODE_BF16) {
    WMat tmp = make_wmat(ctx, &arena, st, A, M, K, fp16);  // pad K to a multiple of 8 and convert, like a weight
    run_gemm(ctx, st, tmp.b16, tmp.ld16, M, w, ep, force_bn);
  } else {
    run_gemm(ctx, st, A, K, M, w, ep, 0);
  }
  F5B_CHECK_CUDA(cudaStreamSynchronize(st));
  F5B_API_END
}

int f5b_op_attention(int32_t mode, const float* qkv, const int32_t* kv_len, int32_t batches, int32_t rows,
                     int32_t heads, float* out, int32_t device, void* stream) {
  F5B_API_BEGIN
  DeviceCtx ctx = open_device(device, mode);
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  const long long M = (long long)batches * rows;
  const int inner = heads * 64;
  if (mode == F5B_MODE_BF16) {
    Arena arena;
    bf16* q16 = arena.alloc_n<bf16>((size_t)M * 3 * inner);
    bf16* o16 = arena.alloc_n<bf16>((size_t)M * inner);
    f32_to_bf16(st, qkv, q16, M * 3 * inner);
    tc_attention(st, q16, o16, batches, rows, heads, kv_len, 0.125f);
    widen_kernel<<<(unsigned)((M * inner + 255) / 256), 256, 0, st>>>(o16, out, M * inner);
    F5B_POST_LAUNCH();
    F5B_CHECK_CUDA(cudaStreamSynchronize(st));
  } else {
    simt_attention(st, qkv, out, batches, rows, heads, kv_len, 0.125f);
    F5B_CHECK_CUDA(cudaStreamSynchronize(st));
  }
  F5B_API_END
}


int f5b_op_attention_bwd(int32_t mode, const float* qkv, const float* d_out, const int32_t* kv_len, int32_t batches,
                         int32_t rows, int32_t heads, float* dqkv, int32_t device, void* stream) {
  F5B_API_BEGIN
  open_device(device, mode);
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  const long long M = (long long)batches * rows;
  const int inner = heads * 64;
  Arena arena;
  float* o = arena.alloc_n<float>((size_t)M * inner);
  float* lse = arena.alloc_n<float>((size_t)M * heads);
  float* D = arena.alloc_n<float>((size_t)M * heads);
  if (mode == F5B_MODE_BF16) {
    bf16* q16 = arena.alloc_n<bf16>((size_t)M * 3 * inner);
    bf16* o16 = arena.alloc_n<bf16>((size_t)M * inner);
    bf16* do16 = arena.alloc_n<bf16>((size_t)M * inner);
    f32_to_bf16(st, qkv, q16, M * 3 * inner);
    tc_attention(st, q16, o16, batches, rows, heads, kv_len, 0.125f, 0, lse);
    widen_kernel<<<(unsigned)((M * inner + 255) / 256), 256, 0, st>>>(o16, o, M * inner);
    F5B_POST_LAUNCH();
    attn_bwd_prep(st, o, d_out, batches, rows, heads, do16, D);
    tc_attention_backward(st, q16, do16, lse, D, batches, rows, heads, kv_len, 0.125f, dqkv);
  } else {
    simt_attention(st, qkv, o, batches, rows, heads, kv_len, 0.125f);
    attention_backward(st, qkv, o, d_out, batches, rows, heads, kv_len, 0.125f, lse, D, dqkv);
  }
  F5B_CHECK_CUDA(cudaStreamSynchronize(st));
  F5B_API_END
}

/* Debug: run the tcgen05 attention once on constant data with the in-kernel timeline enabled; out[3][64][8] receives
 * clock64 stamps of CTA (0,0,0): who 0/1 = softmax tile 0/1 {wait S, S ready, max done, P handed over}, who 2 = MMA
 * issuer {wait P0, P0 seen, wait P1, P1 seen} per key block. */
int f5b_debug_attn_trace(int32_t rows, int32_t heads, int32_t batches, int64_t* out, int32_t device) {
  F5B_API_BEGIN
  open_device(device, F5B_MODE_BF16);
  Arena arena;
  const int inner = heads * 64;
  bf16* Q = arena.alloc_n<bf16>((size_t)batches * rows * 3 * inner);
  bf16* O = arena.alloc_n<bf16>((size_t)batches * rows * inner);
  F5B_CHECK_CUDA(cudaMemset(Q, 0x3c, (size_t)batches * rows * 3 * inner * 2));
  tc_attention(0, Q, O, batches, rows, heads, nullptr, 0.125f);  // warm-up
  F5B_CHECK_CUDA(cudaDeviceSynchronize());
  attn_trace_enable(1);
  tc_attention(0, Q, O, batches, rows, heads, nullptr, 0.125f);
  F5B_CHECK_CUDA(cudaDeviceSynchronize());
  attn_trace_enable(0);
  attn_trace_read(reinterpret_cast<long long*>(out));
  F5B_API_END
}

int f5b_debug_ktrace(int32_t op, int64_t capacity, int64_t* out, int64_t* n, int32_t device) {
  F5B_API_BEGIN
#if defined(F5B_KTRACE)
  static unsigned long long* buf = nullptr;
  static int64_t cap = 0;
  open_device(device, F5B_MODE_BF16);
  auto arm = [](unsigned long long* p) {
    ktrace_set_tc_gemm(p);
    ktrace_set_tc_attn(p);
    ktrace_set_rowwise(p);
  };
  if (op == 0) {
    F5B_REQUIRE(capacity > 0 && buf == nullptr, "trace already running or bad capacity");
    F5B_CHECK_CUDA(cudaMalloc(&buf, (size_t)(4 + 8 * capacity) * 8));
    F5B_CHECK_CUDA(cudaMemset(buf, 0, (size_t)(4 + 8 * capacity) * 8));
    const unsigned long long c = (unsigned long long)capacity;
    F5B_CHECK_CUDA(cudaMemcpy(buf + 1, &c, 8, cudaMemcpyHostToDevice));
    cap = capacity;
    arm(buf);
  } else {
    F5B_REQUIRE(buf != nullptr && out != nullptr && n != nullptr, "trace not running");
    F5B_CHECK_CUDA(cudaDeviceSynchronize());
    arm(nullptr);
    unsigned long long cnt = 0;
    F5B_CHECK_CUDA(cudaMemcpy(&cnt, buf, 8, cudaMemcpyDeviceToHost));
    const int64_t take = std::min<int64_t>(std::min<int64_t>((int64_t)cnt, cap), *n);
    F5B_CHECK_CUDA(cudaMemcpy(out, buf + 4, (size_t)take * 64, cudaMemcpyDeviceToHost));
    *n = (int64_t)cnt;
    cudaFree(buf);
    buf = nullptr;
  }
#else
  (void)op; (void)capacity; (void)out; (void)n; (void)device;
  F5B_REQUIRE(false, "library built without F5B_KTRACE");
#endif
  F5B_API_END
}

int f5b_op_time(int32_t which, int32_t M, int32_t N, int32_t K, int32_t iters, float* ms, int32_t device, void* stream) {
  F5B_API_BEGIN
  F5B_REQUIRE(ms != nullptr && iters > 0, "bad argument");
  DeviceCtx ctx = open_device(device, F5B_MODE_BF16);
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  Arena arena;
  cudaEvent_t e0, e1;
  F5B_CHECK_CUDA(cudaEventCreate(&e0));
  F5B_CHECK_CUDA(cudaEventCreate(&e1));
  // `iters` launches are captured into one CUDA graph and the graph is timed: host launch cost (three tensor-map
  // encodes per GEMM) would otherwise bound kernels of a few microseconds
  cudaStream_t cs = nullptr;
  F5B_CHECK_CUDA(cudaStreamCreateWithFlags(&cs, cudaStreamNonBlocking));
  cudaStream_t user_st = st;
  st = cs;
  auto timed = [&](auto&& fn) {
    fn();
    F5B_CHECK_CUDA(cudaStreamSynchronize(cs));
    cudaGraph_t g = nullptr;
    cudaGraphExec_t ge = nullptr;
    long long c0 = launch_count();
    F5B_CHECK_CUDA(cudaStreamBeginCapture(cs, cudaStreamCaptureModeThreadLocal));
    for (int i = 0; i < iters; ++i) fn();
    F5B_CHECK_CUDA(cudaStreamEndCapture(cs, &g));
    count_launch(-(launch_count() - c0));
    F5B_CHECK_CUDA(cudaGraphInstantiate(&ge, g, 0));
    F5B_CHECK_CUDA(cudaStreamSynchronize(user_st));
    F5B_CHECK_CUDA(cudaGraphLaunch(ge, cs));  // warm-up replay
    F5B_CHECK_CUDA(cudaEventRecord(e0, cs));
    F5B_CHECK_CUDA(cudaGraphLaunch(ge, cs));
    F5B_CHECK_CUDA(cudaEventRecord(e1, cs));
    F5B_CHECK_CUDA(cudaEventSynchronize(e1));
    count_launch(2LL * iters);
    float t = 0;
    F5B_CHECK_CUDA(cudaEventElapsedTime(&t, e0, e1));
    *ms = t / iters;
    cudaGraphExecDestroy(ge);
    cudaGraphDestroy(g);
  };
  if (which == 0 || which == 4 || which == 5 || which == 6 || which == 64 || which == 128 || which == 192 || which == 256 || which > 1000) {
    // 0: bf16 store, tile width chosen by the library; 64/128/256: forced tile width; 4: gate*y + residual (fp32 reduce)
    long long ld = (K + 7) / 8 * 8;
    bf16* A = arena.alloc_n<bf16>((size_t)M * ld);
    // in the sampling loop every GEMM streams its weights from HBM (0.66 GB per evaluation > 126 MB L2): rotate over
    // enough weight copies that none is L2-resident when its turn comes again
    const size_t wbytes = (size_t)N * ld * 2;
    int ncopy = (int)((size_t)(192u << 20) / wbytes) + 1;
    if (ncopy > iters) ncopy = iters;
    bf16* W = arena.alloc_n<bf16>((size_t)N * ld * ncopy);
    bf16* O = arena.alloc_n<bf16>((size_t)M * N);
    fill_random16(st, A, (long long)M * ld, 1.0f, 1u);                          // LN-like activations
    fill_random16(st, W, (long long)N * ld * ncopy, 1.7f / sqrtf((float)K), 2u);  // std ~ 1 / sqrt(K)
    EpiParams ep;
    ep.out = O;
    ep.out_type = 1;
    ep.ldo = N;
    ep.n_valid = N;
    if (which == 4 || which == 5 || which == 6) {  // 5 / 6: residual epilogue, 256-wide tiles, split-K 2 / 128-wide, split-K 2
      float* R = arena.alloc_n<float>((size_t)M * N);
      F5B_CHECK_CUDA(cudaMemsetAsync(R, 0, (size_t)M * N * 4, st));
      ep.kind = EPI_RESID;
      ep.resid = R;
      ep.ldr = N;
    }
    const int fbn = which >= 64 ? which : which == 5 ? 256 : which == 6 ? 128 : 0;
    const int fsp = which == 5 || which == 6 ? 2 : 0;
    int turn = 0;
    timed([&] {
      tc_gemm(st, ctx.sm_count, A, ld, M, 1, W + (size_t)(turn % ncopy) * N * ld, ld, N, K, ep, nullptr, fbn, fsp);
      ++turn;
    });
  } else if (which == 1) {
    const int rows = M, heads = N, batches = K, inner = heads * 64;
    bf16* Q = arena.alloc_n<bf16>((size_t)batches * rows * 3 * inner);
    bf16* O = arena.alloc_n<bf16>((size_t)batches * rows * inner);
    fill_random16(st, Q, (long long)batches * rows * 3 * inner, 1.7f, 3u);  // unit-variance q, k, v: scores ~ N(0, 1)
    timed([&] { tc_attention(st, Q, O, batches, rows, heads, nullptr, 0.125f); });
  } else if (which == 2) {
    float* X = arena.alloc_n<float>((size_t)M * N);
    bf16* O = arena.alloc_n<bf16>((size_t)M * N);
    float* V = arena.alloc_n<float>((size_t)2 * N);
    fill_random32_kernel<<<(unsigned)(((long long)M * N + 255) / 256), 256, 0, st>>>(X, (long long)M * N, 1.7f, 4u);
    F5B_POST_LAUNCH();
    fill_random32_kernel<<<(unsigned)((2 * N + 255) / 256), 256, 0, st>>>(V, 2 * N, 0.1f, 5u);
    F5B_POST_LAUNCH();
    timed([&] { ln_mod<bf16>(st, X, O, M, N, V, V + N, 1.0f, 0, 1 << 30, 1, 1e-6f); });
  } else {
    F5B_REQUIRE(false, "unknown kernel selector");
  }
  cudaEventDestroy(e0);
  cudaEventDestroy(e1);
  cudaStreamDestroy(cs);
  F5B_API_END
}

}  // extern "C"
