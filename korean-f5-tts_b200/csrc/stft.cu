// Mel-filterbank STFT frontend and inverse STFT / overlap-add, hand-written (no cuFFT).
//
//   mel:   reference src/f5_tts/model/modules.py:78-104 (torchaudio MelSpectrogram: reflect pad 512, periodic Hann 1024,
//          hop 256, |rFFT|, HTK filterbank 513->100 f in [0, 12000], clamp(1e-5).log()).
//   istft: Vocos ISTFTHead (third-party vocos; witnesses in the reference: runtime/triton_trtllm/scripts/
//          export_vocoder_to_onnx.py:51-59 and scripts/conv_stft.py:193-234): exp, clip(1e2), cos/sin, irFFT 1024,
//          Hann window, overlap-add, divide by the window-square envelope, trim 512 each side.
// One CTA per frame; the 1024-point transform is a radix-2 Stockham-free in-place FFT in shared memory with an fp32
// twiddle table generated in double precision.
#include <cmath>
#include <mutex>
#include <vector>

#include "plan.h"

namespace f5b {

static constexpr int NFFT = 1024, HOP = 256, NBIN = 513, NMEL = 100;

struct StftTables {
  float2* tw = nullptr;   // [512] (cos, sin)(2 pi k / 1024)
  float* win = nullptr;   // [1024] periodic Hann
  float* fb = nullptr;    // [513, 100]
  int device = -1;
};
static StftTables g_tabs[64];  // one set per device, built on first use with that device current
static std::mutex g_tab_mu;

static const StftTables& tables_for(int device) {
  std::lock_guard<std::mutex> lk(g_tab_mu);
  F5B_REQUIRE(device >= 0 && device < 64, "device index beyond the STFT table slots");
  StftTables& g_tab = g_tabs[device];
  if (g_tab.device == device) return g_tab;
  F5B_CHECK_CUDA(cudaSetDevice(device));
  std::vector<float2> tw(512);
  std::vector<float> win(NFFT), fb((size_t)NBIN * NMEL);
  const double PI = 3.14159265358979323846;
  for (int k = 0; k < 512; ++k) tw[k] = make_float2((float)std::cos(2 * PI * k / NFFT), (float)std::sin(2 * PI * k / NFFT));
  for (int n = 0; n < NFFT; ++n) win[n] = (float)(0.5 - 0.5 * std::cos(2 * PI * n / NFFT));
  // torchaudio.functional.melscale_fbanks(513, 0, 12000, 100, 24000, norm=None, mel_scale="htk")
  const double m_min = 2595.0 * std::log10(1.0 + 0.0 / 700.0), m_max = 2595.0 * std::log10(1.0 + 12000.0 / 700.0);
  std::vector<double> f_pts(NMEL + 2);
  for (int i = 0; i < NMEL + 2; ++i) {
    double m = m_min + (m_max - m_min) * i / (NMEL + 1);
    f_pts[i] = 700.0 * (std::pow(10.0, m / 2595.0) - 1.0);
  }
  for (int k = 0; k < NBIN; ++k) {
    double f = 12000.0 * k / (NBIN - 1);
    for (int c = 0; c < NMEL; ++c) {
      double down = (f - f_pts[c]) / (f_pts[c + 1] - f_pts[c]);
      double up = (f_pts[c + 2] - f) / (f_pts[c + 2] - f_pts[c + 1]);
      double v = down < up ? down : up;
      fb[(size_t)k * NMEL + c] = (float)(v > 0 ? v : 0);
    }
  }
  F5B_CHECK_CUDA(cudaMalloc(&g_tab.tw, 512 * sizeof(float2)));
  F5B_CHECK_CUDA(cudaMalloc(&g_tab.win, NFFT * 4));
  F5B_CHECK_CUDA(cudaMalloc(&g_tab.fb, fb.size() * 4));
  F5B_CHECK_CUDA(cudaMemcpy(g_tab.tw, tw.data(), 512 * sizeof(float2), cudaMemcpyHostToDevice));
  F5B_CHECK_CUDA(cudaMemcpy(g_tab.win, win.data(), NFFT * 4, cudaMemcpyHostToDevice));
  F5B_CHECK_CUDA(cudaMemcpy(g_tab.fb, fb.data(), fb.size() * 4, cudaMemcpyHostToDevice));
  g_tab.device = device;
  return g_tab;
}

// In-place 1024-point complex FFT over shared arrays (bit-reversed input order expected), 256 threads.
// sign = -1: forward (e^{-i...}), +1: inverse (unnormalised).
__device__ __forceinline__ void fft1024_smem(float* re, float* im, const float2* __restrict__ tw, float sign) {
  for (int s = 1; s <= 10; ++s) {
    const int half = 1 << (s - 1);
    __syncthreads();
#pragma unroll
    for (int r = 0; r < 2; ++r) {
      const int j = threadIdx.x + r * 256;  // butterfly index 0..511
      const int grp = j >> (s - 1), pos = j & (half - 1);
      const int i0 = (grp << s) + pos, i1 = i0 + half;
      const float2 w = tw[pos << (10 - s)];
      const float wr = w.x, wi = sign * w.y;
      const float xr = re[i1], xi = im[i1];
      const float tr = xr * wr - xi * wi, ti = xr * wi + xi * wr;
      const float ur = re[i0], ui = im[i0];
      re[i0] = ur + tr; im[i0] = ui + ti;
      re[i1] = ur - tr; im[i1] = ui - ti;
    }
  }
  __syncthreads();
}

__global__ void __launch_bounds__(256)
mel_kernel(const float* __restrict__ wav, int nw, int T, const float2* __restrict__ tw, const float* __restrict__ win,
           const float* __restrict__ fb, float* __restrict__ mel) {
  __shared__ float re[NFFT], im[NFFT];
  __shared__ float mag[NBIN + 3];
  const int t = blockIdx.x, b = blockIdx.y;
  const float* w = wav + (long long)b * nw;
  for (int n = threadIdx.x; n < NFFT; n += 256) {
    int idx = t * HOP - NFFT / 2 + n;
    if (idx < 0) idx = -idx;
    if (idx >= nw) idx = 2 * (nw - 1) - idx;
    const int br = __brev((unsigned)n) >> 22;
    re[br] = w[idx] * win[n];
    im[br] = 0.0f;
  }
  fft1024_smem(re, im, tw, -1.0f);
  for (int k = threadIdx.x; k < NBIN; k += 256) mag[k] = sqrtf(re[k] * re[k] + im[k] * im[k]);
  __syncthreads();
  if (threadIdx.x < NMEL) {
    const int c = threadIdx.x;
    float acc = 0.0f;
    for (int k = 0; k < NBIN; ++k) acc = fmaf(mag[k], __ldg(fb + k * NMEL + c), acc);
    mel[((long long)b * NMEL + c) * T + t] = logf(fmaxf(acc, 1e-5f));
  }
}

// Vocos ISTFTHead after the linear layer, fused: exp / clip / cos / sin, inverse real FFT, Hann window, overlap-add,
// division by the window-square envelope, trim.  One CTA owns 1024 consecutive OUTPUT samples (4 hops) of one
// utterance and transforms the 7 frames that overlap them itself (1.75x redundant FFT work -- the transform is ~50 kflop)
// so the windowed frames never travel through HBM: the unfused pair wrote [T, 1024] fp32 frames and re-read them 4x,
// ~5 KB of traffic per frame against 4.1 KB of spectrum in + 1 KB of samples out here.
//   z [B * Tmax, ldz]: cols [0,513) log-magnitude, [513,1026) phase;  lens[b] (or Tmax) frames are valid;
//   out [B, (Tmax - 1) * 256]: row b holds (lens[b] - 1) * 256 samples, zeros after them.
__global__ void __launch_bounds__(256)
istft_ola_kernel(const float* __restrict__ z, long long ldz, const float2* __restrict__ tw,
                 const float* __restrict__ win, const int* __restrict__ lens, int Tmax, float* __restrict__ out) {
  __shared__ float re[NFFT], im[NFFT], acc[NFFT];
  const int b = blockIdx.y, c = blockIdx.x;
  const int T = lens ? lens[b] : Tmax;
  const int L = (T - 1) * HOP, Lmax = (Tmax - 1) * HOP;
  const int s0 = c * NFFT;
  float* o = out + (long long)b * Lmax;
  if (s0 >= L) {  // beyond this utterance: defined output (zeros)
    for (int s = threadIdx.x; s < NFFT; s += 256)
      if (s0 + s < Lmax) o[s0 + s] = 0.0f;
    return;
  }
  for (int s = threadIdx.x; s < NFFT; s += 256) acc[s] = 0.0f;
  for (int t = 4 * c - 1; t <= 4 * c + 5; ++t) {
    if (t < 0 || t >= T) continue;  // uniform
    const float* zr = z + ((long long)b * Tmax + t) * ldz;
    __syncthreads();  // previous frame's readers of re[] are done
    for (int k = threadIdx.x; k < NBIN; k += 256) {
      const float mg = fminf(expf(zr[k]), 100.0f);
      float sn, cs;
      sincosf(zr[NBIN + k], &sn, &cs);
      float xr = mg * cs, xi = mg * sn;
      if (k == 0 || k == NFFT / 2) xi = 0.0f;  // irfft ignores the imaginary part of DC / Nyquist
      const int br = __brev((unsigned)k) >> 22;
      re[br] = xr;
      im[br] = xi;
      if (k > 0 && k < NFFT / 2) {
        const int br2 = __brev((unsigned)(NFFT - k)) >> 22;
        re[br2] = xr;
        im[br2] = -xi;
      }
    }
    fft1024_smem(re, im, tw, 1.0f);
    for (int n = threadIdx.x; n < NFFT; n += 256) {
      const int s = t * HOP + n - NFFT / 2 - s0;  // distinct n -> distinct s: no conflicts
      if (s >= 0 && s < NFFT) acc[s] += re[n] * (1.0f / NFFT) * win[n];
    }
  }
  __syncthreads();
  for (int s = threadIdx.x; s < NFFT; s += 256) {
    if (s0 + s >= Lmax) continue;
    float v = 0.0f;
    if (s0 + s < L) {
      const int pos = s0 + s + NFFT / 2;
      int t_hi = pos / HOP;
      if (t_hi > T - 1) t_hi = T - 1;
      float env = 0.0f;
      for (int t = t_hi; t >= 0 && pos - t * HOP < NFFT; --t) {
        const float w = win[pos - t * HOP];
        env = fmaf(w, w, env);
      }
      v = acc[s] / env;
    }
    o[s0 + s] = v;
  }
}

void istft_run(cudaStream_t st, int device, const float* z, long long ldz, int B, int T, const int* lens, float* wav) {
  const StftTables& tb = tables_for(device);
  const int Lmax = (T - 1) * HOP;
  if (Lmax <= 0) return;
  istft_ola_kernel<<<dim3((Lmax + NFFT - 1) / NFFT, B), 256, 0, st>>>(z, ldz, tb.tw, tb.win, lens, T, wav);
  F5B_POST_LAUNCH();
}

void stft_prepare(int device) { tables_for(device); }

DeviceCtx open_device(int device, int mode);

}  // namespace f5b

using namespace f5b;

extern "C" int f5b_mel(const float* wav, int32_t B, int32_t nw, float* mel, int32_t device, int32_t mem, void* stream) {
  try {
    F5B_REQUIRE(wav && mel && B > 0, "null argument");
    F5B_REQUIRE(nw > NFFT / 2, "waveform shorter than the reflect padding (torch.stft raises too)");
    open_device(device, F5B_MODE_FP32);
    cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
    const StftTables& tb = tables_for(device);
    const int T = nw / HOP + 1;
    Arena arena;
    const float* dw = wav;
    float* dm = mel;
    if (mem == F5B_MEM_HOST) {
      float* w = arena.alloc_n<float>((size_t)B * nw);
      copy_in(w, wav, (size_t)B * nw * 4, mem, st);
      dw = w;
      dm = arena.alloc_n<float>((size_t)B * NMEL * T);
    }
    long long c0 = launch_count();
    mel_kernel<<<dim3(T, B), 256, 0, st>>>(dw, nw, T, tb.tw, tb.win, tb.fb, dm);
    F5B_POST_LAUNCH();
    set_last_call(launch_count() - c0);
    if (mem == F5B_MEM_HOST) copy_out(mel, dm, (size_t)B * NMEL * T * 4, mem, st);
    F5B_CHECK_CUDA(cudaStreamSynchronize(st));
  } catch (const f5b::Err& e) {
    return e.code;
  }
  return F5B_OK;
}
