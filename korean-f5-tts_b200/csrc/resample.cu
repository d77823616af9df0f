// Reference-audio front half (reference infer/utils_infer.py:523-533): channel mix-down, RMS
// measurement / lift to the target level, and the windowed-sinc polyphase resampler to 24 kHz
// (torchaudio.transforms.Resample: a strided FIR, one filter row per output phase).  Memory-bound: every input sample
// is read from HBM once (staged in shared memory per block of output frames), every output sample written once.
#include <cmath>
#include <vector>

#include "plan.h"

namespace f5b {
DeviceCtx open_device(int device, int mode);
namespace {

constexpr int RS_THREADS = 256;
constexpr int RS_PART = 1024;  // partial sums of the RMS reduction (fixed tree => deterministic)

__device__ __forceinline__ float mono_at(const float* __restrict__ wav, int channels, long long nw, long long i) {
  float s = 0.f;
  for (int c = 0; c < channels; ++c) s += __ldg(wav + (size_t)c * nw + i);
  return channels > 1 ? s / (float)channels : s;
}

// partial[b] = sum over this block's slice of mono(x)^2
__global__ void __launch_bounds__(RS_THREADS) mono_sumsq_kernel(const float* __restrict__ wav, int channels,
                                                                long long nw, float* __restrict__ partial) {
  __shared__ float red[RS_THREADS / 32];
  const long long per = (nw + gridDim.x - 1) / gridDim.x;
  const long long lo = per * blockIdx.x, hi = lo + per < nw ? lo + per : nw;
  float acc = 0.f;
  for (long long i = lo + threadIdx.x; i < hi; i += RS_THREADS) {
    const float v = mono_at(wav, channels, nw, i);
    acc = fmaf(v, v, acc);
  }
  acc = warp_sum(acc);
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = acc;
  __syncthreads();
  if (threadIdx.x < 32) {
    float v = threadIdx.x < RS_THREADS / 32 ? red[threadIdx.x] : 0.f;
    v = warp_sum(v);
    if (threadIdx.x == 0) partial[blockIdx.x] = v;
  }
}

// stats[0] = rms, stats[1] = gain (target / rms when rms < target, else 1; utils_infer.py:527-529)
__global__ void __launch_bounds__(RS_THREADS) rms_finish_kernel(const float* __restrict__ partial, int n_part,
                                                                long long nw, float target_rms,
                                                                float* __restrict__ stats) {
  __shared__ float red[RS_THREADS / 32];
  float acc = 0.f;
  for (int i = threadIdx.x; i < n_part; i += RS_THREADS) acc += partial[i];
  acc = warp_sum(acc);
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = acc;
  __syncthreads();
  if (threadIdx.x == 0) {
    float s = 0.f;
    for (int w = 0; w < RS_THREADS / 32; ++w) s += red[w];
    const float rms = sqrtf(s / (float)nw);
    stats[0] = rms;
    stats[1] = (target_rms > 0.f && rms < target_rms) ? target_rms / rms : 1.f;
  }
}

// out[f * n_new + j] = gain * sum_k table[j][k] * xpad[f * n_orig + k], xpad = mono(x) zero-padded by `width` in front.
// One block = `fb` consecutive output frames; its input span ((fb-1) * n_orig + taps samples) is staged in smem.
__global__ void __launch_bounds__(RS_THREADS) resample_kernel(const float* __restrict__ wav, int channels,
                                                              long long nw, const float* __restrict__ table,
                                                              int n_new, int taps, int n_orig, int width, int fb,
                                                              const float* __restrict__ stats,
                                                              float* __restrict__ out, long long out_len) {
  extern __shared__ float span[];
  const long long f0 = (long long)blockIdx.x * fb;
  const long long in0 = f0 * n_orig - width;
  const int span_n = (fb - 1) * n_orig + taps;
  const float gain = stats[1];
  for (int s = threadIdx.x; s < span_n; s += RS_THREADS) {
    const long long p = in0 + s;
    span[s] = (p >= 0 && p < nw) ? mono_at(wav, channels, nw, p) * gain : 0.f;
  }
  __syncthreads();
  const int n_out = fb * n_new;
  for (int o = threadIdx.x; o < n_out; o += RS_THREADS) {
    const long long i = f0 * n_new + o;
    if (i >= out_len) break;
    const int f = o / n_new, j = o - f * n_new;
    const float* __restrict__ h = table + (size_t)j * taps;
    const float* x = span + f * n_orig;
    float acc = 0.f;
    for (int k = 0; k < taps; ++k) acc = fmaf(__ldg(h + k), x[k], acc);
    out[i] = acc;
  }
}

}  // namespace
}  // namespace f5b

using namespace f5b;

extern "C" int f5b_ref_audio(const float* wav, int32_t channels, int64_t nw, const float* table, int32_t n_new,
                             int32_t taps, int32_t n_orig, int32_t width, float target_rms, float* out,
                             int64_t out_len, float* rms_host, int32_t device, int32_t mem, void* stream) {
  try {
    F5B_REQUIRE(wav && table && out, "null argument");
    F5B_REQUIRE(channels > 0 && nw > 0 && out_len > 0, "empty waveform");
    F5B_REQUIRE(n_new > 0 && n_orig > 0 && taps > 0 && width >= 0 && taps >= n_orig, "bad resampling filter geometry");
    F5B_REQUIRE(out_len <= ((nw + n_orig - 1) / n_orig + 1) * (int64_t)n_new, "out_len exceeds the resampled length");
    F5B_REQUIRE((int64_t)n_new * taps <= (64ll << 20), "resampling filter bank too large (sample rates nearly coprime)");
    open_device(device, F5B_MODE_FP32);
    cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
    int fb = 512 / n_new;
    if (fb < 1) fb = 1;
    while (fb > 1 && (size_t)((fb - 1) * (int64_t)n_orig + taps) * 4 > 40960) fb >>= 1;
    const size_t smem = (size_t)((fb - 1) * (int64_t)n_orig + taps) * 4;
    F5B_REQUIRE(smem <= 200 * 1024, "resampling filter longer than the shared-memory stage");
    if (smem > 48 * 1024)
      F5B_CHECK_CUDA(cudaFuncSetAttribute(resample_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));

    Arena arena;
    const float *dw = wav, *dt = table;
    float* dout = out;
    if (mem == F5B_MEM_HOST) {
      float* w = arena.alloc_n<float>((size_t)channels * nw);
      copy_in(w, wav, (size_t)channels * nw * 4, mem, st);
      float* t = arena.alloc_n<float>((size_t)n_new * taps);
      copy_in(t, table, (size_t)n_new * taps * 4, mem, st);
      dw = w;
      dt = t;
      dout = arena.alloc_n<float>((size_t)out_len);
    }
    float* partial = arena.alloc_n<float>(RS_PART + 2);
    float* stats = partial + RS_PART;
    int n_part = (int)((nw + 4095) / 4096);
    if (n_part > RS_PART) n_part = RS_PART;
    long long c0 = launch_count();
    mono_sumsq_kernel<<<n_part, RS_THREADS, 0, st>>>(dw, channels, nw, partial);
    F5B_POST_LAUNCH();
    rms_finish_kernel<<<1, RS_THREADS, 0, st>>>(partial, n_part, nw, target_rms, stats);
    F5B_POST_LAUNCH();
    const long long frames = (out_len + n_new - 1) / n_new;
    const long long blocks = (frames + fb - 1) / fb;
    resample_kernel<<<(unsigned)blocks, RS_THREADS, smem, st>>>(dw, channels, nw, dt, n_new, taps, n_orig, width, fb,
                                                                stats, dout, out_len);
    F5B_POST_LAUNCH();
    set_last_call(launch_count() - c0);
    if (mem == F5B_MEM_HOST) copy_out(out, dout, (size_t)out_len * 4, mem, st);
    float rms = 0.f;
    F5B_CHECK_CUDA(cudaMemcpyAsync(&rms, stats, 4, cudaMemcpyDeviceToHost, st));
    F5B_CHECK_CUDA(cudaStreamSynchronize(st));
    if (rms_host) *rms_host = rms;
  } catch (const f5b::Err& e) {
    return e.code;
  }
  return F5B_OK;
}
