// f5b_vocos: Vocos mel -> waveform (third-party `vocos` package, charactr/vocos-mel-24khz hyper-parameters; call
// sites in the reference: infer/utils_infer.py:114-137,703, model/cfm.py:225-227).
//   Conv1d(100->512, k7) as im2col + GEMM -> LN -> 8 x [dwconv k7 + LN | pwconv1 + GELU | pwconv2 * gamma + residual]
//   -> LN -> Linear(512 -> 1026) -> exp/clip/cos/sin -> iSTFT (stft.cu).
#include <mutex>

#include "plan.h"

namespace f5b {

DeviceCtx open_device(int device, int mode);
void istft_run(cudaStream_t st, int device, const float* z, long long ldz, int B, int T, const int* lens, float* wav);
void stft_prepare(int device);

static constexpr int V_MEL = 100, V_DIM = 512, V_INT = 1536, V_LAYERS = 8, V_OUT = 1026, V_OUT_LD = 1028;
static constexpr int V_KIN = V_MEL * 7, V_KIN_LD = 704;

// mel [B, 100, T] -> A [B*T, 704]: A[m, ci*7 + k] = mel[b, ci, t + k - 3]
template <typename T>
__global__ void __launch_bounds__(256)
vocos_im2col_kernel(const float* __restrict__ mel, T* __restrict__ out, int Tn, const int* __restrict__ lens,
                    long long total) {
  const long long i = (long long)blockIdx.x * 256 + threadIdx.x;
  if (i >= total) return;
  const long long m = i / V_KIN_LD;
  const int c = (int)(i - m * V_KIN_LD);
  const int b = (int)(m / Tn), t = (int)(m - (long long)b * Tn);
  float v = 0.0f;
  if (c < V_KIN) {
    const int ci = c / 7, k = c - ci * 7;
    const int tt = t + k - 3;
    if (tt >= 0 && tt < (lens ? lens[b] : Tn)) v = mel[((long long)b * V_MEL + ci) * Tn + tt];  // zero padding at the utterance's own end
  }
  st_from_float(out + i, v);
}

struct VBlock {
  const float *dw_t, *dw_b, *ln_w, *ln_b, *b1, *b2, *gamma;
  WMat pw1, pw2;
};

}  // namespace f5b

using namespace f5b;

namespace f5b { __global__ void dw_transpose(const float* __restrict__ w, float* __restrict__ out, int d); }

struct f5b_vocos {
  DeviceCtx ctx;
  Arena warena;
  std::unique_ptr<WeightStore> ws;
  std::unique_ptr<Arena> wk;
  long long cap_rows = 0;
  WMat embed, head;
  const float *embed_b, *norm_w, *norm_b, *fin_w, *fin_b, *head_b;
  std::vector<VBlock> blocks;
  float *mel_in, *x0, *x, *z, *wav;
  int* lens_dev;
  void *a_in, *norm, *h;
  cudaEvent_t done = nullptr;  // end of the last decode: the next call (any stream) orders itself after it
  std::mutex mu;

  f5b_vocos(const f5b_tensor* w, int n, int mode, int device) {
    ctx = open_device(device, mode);
    stft_prepare(device);
    cudaStream_t st = 0;
    ws.reset(new WeightStore(&warena, w, n, st));
    auto W = [&](const std::string& nm, long long numel) { return ws->get(nm, numel); };
    embed = make_wmat(ctx, &warena, st, W("backbone.embed.weight", (long long)V_DIM * V_KIN), V_DIM, V_KIN);
    embed_b = W("backbone.embed.bias", V_DIM);
    norm_w = W("backbone.norm.weight", V_DIM);
    norm_b = W("backbone.norm.bias", V_DIM);
    for (int i = 0; i < V_LAYERS; ++i) {
      std::string p = "backbone.convnext." + std::to_string(i);
      VBlock b;
      float* dwt = warena.alloc_n<float>(7 * V_DIM);
      dw_transpose<<<(V_DIM * 7 + 255) / 256, 256, 0, st>>>(W(p + ".dwconv.weight", V_DIM * 7), dwt, V_DIM);
      F5B_POST_LAUNCH();
      b.dw_t = dwt;
      b.dw_b = W(p + ".dwconv.bias", V_DIM);
      b.ln_w = W(p + ".norm.weight", V_DIM);
      b.ln_b = W(p + ".norm.bias", V_DIM);
      b.pw1 = make_wmat(ctx, &warena, st, W(p + ".pwconv1.weight", (long long)V_INT * V_DIM), V_INT, V_DIM);
      b.b1 = W(p + ".pwconv1.bias", V_INT);
      b.pw2 = make_wmat(ctx, &warena, st, W(p + ".pwconv2.weight", (long long)V_INT * V_DIM), V_DIM, V_INT);
      b.b2 = W(p + ".pwconv2.bias", V_DIM);
      b.gamma = W(p + ".gamma", V_DIM);
      blocks.push_back(b);
    }
    fin_w = W("backbone.final_layer_norm.weight", V_DIM);
    fin_b = W("backbone.final_layer_norm.bias", V_DIM);
    head = make_wmat(ctx, &warena, st, W("head.out.weight", (long long)V_OUT * V_DIM), V_OUT, V_DIM);
    head_b = W("head.out.bias", V_OUT);
    F5B_CHECK_CUDA(cudaStreamSynchronize(st));
    F5B_CHECK_CUDA(cudaEventCreateWithFlags(&done, cudaEventDisableTiming));
    lens_dev = warena.alloc_n<int>(kMaxBatch);
  }
  ~f5b_vocos() {
    if (done) cudaEventDestroy(done);
  }

  static constexpr int kMaxBatch = 1 << 16;  // lens_dev is allocated once: a growing batch size must not re-allocate

  void ensure(long long rows, int B) {
    F5B_REQUIRE(B <= kMaxBatch, "Vocos batch larger than 65536 utterances");
    if (wk && rows <= cap_rows) return;
    F5B_CHECK_CUDA(cudaDeviceSynchronize());
    const size_t es = ctx.mode == F5B_MODE_BF16 ? 2 : 4;
    rows += rows / 8;  // headroom: slightly larger requests do not re-allocate
    wk.reset();
    cap_rows = 0;
    std::unique_ptr<Arena> fresh(new Arena());
    Arena& A = *fresh;
    mel_in = A.alloc_n<float>((size_t)rows * V_MEL);
    x0 = A.alloc_n<float>((size_t)rows * V_DIM);
    x = A.alloc_n<float>((size_t)rows * V_DIM);
    z = A.alloc_n<float>((size_t)rows * V_OUT_LD);
    wav = A.alloc_n<float>((size_t)rows * 256);
    a_in = A.alloc((size_t)rows * V_KIN_LD * es);
    norm = A.alloc((size_t)rows * V_DIM * es);
    h = A.alloc((size_t)rows * V_INT * es);
    wk = std::move(fresh);  // commit: a failed growth leaves no workspace and zero capacity behind
    cap_rows = rows;
  }

  template <typename T>
  void run(cudaStream_t st, int B, int Tn, const int* lens) {
    const long long M = (long long)B * Tn;
    const bool is16 = sizeof(T) == 2;
    const long long total = M * V_KIN_LD;
    vocos_im2col_kernel<T><<<(unsigned)((total + 255) / 256), 256, 0, st>>>(mel_in, (T*)a_in, Tn, lens, total);
    F5B_POST_LAUNCH();
    EpiParams e;
    e.bias = embed_b;
    e.out = x0;
    e.ldo = V_DIM;
    e.n_valid = V_DIM;
    run_gemm(ctx, st, a_in, V_KIN_LD, M, embed, e);
    ln_mod<float>(st, x0, x, M, V_DIM, norm_w, norm_b, 0.0f, 0, 1 << 30, 1, 1e-6f);
    for (const VBlock& b : blocks) {
      dwconv7_ln<T>(st, x, (T*)norm, B, Tn, V_DIM, lens, b.dw_t, b.dw_b, b.ln_w, b.ln_b, 1e-6f);
      EpiParams e1;
      e1.act = ACT_GELU_ERF;
      e1.bias = b.b1;
      e1.out = h;
      e1.out_type = is16 ? OUT_F16 : OUT_F32;
      e1.ldo = V_INT;
      e1.n_valid = V_INT;
      run_gemm(ctx, st, norm, V_DIM, M, b.pw1, e1);
      EpiParams e2;
      e2.kind = EPI_RESID;
      e2.bias = b.b2;
      e2.resid = x;
      e2.ldr = V_DIM;
      e2.n_valid = V_DIM;
      e2.gate = b.gamma;  // layer scale
      run_gemm(ctx, st, h, V_INT, M, b.pw2, e2);
    }
    ln_mod<T>(st, x, (T*)norm, M, V_DIM, fin_w, fin_b, 0.0f, 0, 1 << 30, 1, 1e-6f);
    EpiParams eh;
    eh.bias = head_b;
    eh.out = z;
    eh.ldo = V_OUT_LD;
    eh.n_valid = V_OUT;
    run_gemm(ctx, st, norm, V_DIM, M, head, eh);
    istft_run(st, ctx.device, z, V_OUT_LD, B, Tn, lens, wav);
  }
};

extern "C" {

int f5b_vocos_create(const f5b_tensor* weights, int32_t n_weights, int32_t mode, int32_t device, f5b_vocos** out) {
  try {
    F5B_REQUIRE(weights && out && n_weights > 0, "null argument");
    F5B_REQUIRE(mode == F5B_MODE_FP32 || mode == F5B_MODE_BF16, "unknown precision mode");
    *out = nullptr;
    std::unique_ptr<f5b_vocos> v(new f5b_vocos(weights, n_weights, mode, device));
    *out = v.release();
  } catch (const f5b::Err& e) {
    return e.code;
  }
  return F5B_OK;
}

void f5b_vocos_destroy(f5b_vocos* v) {
  if (!v) return;
  cudaSetDevice(v->ctx.device);
  cudaDeviceSynchronize();
  delete v;
}

// mel [B, 100, T] (+ lens [B] int32 or NULL: row b holds lens[b] <= T frames, zero padding applies at ITS end) ->
// wav [B, (T - 1) * 256], row b valid up to (lens[b] - 1) * 256 samples, zeros after.  Device buffers: asynchronous.
int f5b_vocos_reserve(f5b_vocos* v, int32_t B, int32_t T) {
  try {
    F5B_REQUIRE(v && B > 0 && T > 1, "bad argument");
    F5B_CHECK_CUDA(cudaSetDevice(v->ctx.device));
    std::lock_guard<std::mutex> lock(v->mu);
    v->ensure((long long)B * T, B);
  } catch (const f5b::Err& e) {
    return e.code;
  }
  return F5B_OK;
}

int f5b_vocos_decode_varlen(f5b_vocos* v, const float* mel, const int32_t* lens, int32_t B, int32_t T, float* wav,
                            int32_t mem, void* stream) {
  try {
    F5B_REQUIRE(v && mel && wav, "null argument");
    F5B_REQUIRE(B > 0 && T > 1, "Vocos.decode needs at least 2 frames");
    F5B_CHECK_CUDA(cudaSetDevice(v->ctx.device));
    std::lock_guard<std::mutex> lock(v->mu);
    cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
    F5B_CHECK_CUDA(cudaStreamWaitEvent(st, v->done, 0));
    const long long M = (long long)B * T;
    v->ensure(M, B);
    copy_in(v->mel_in, mel, (size_t)M * V_MEL * 4, mem, st);
    if (lens) copy_in(v->lens_dev, lens, (size_t)B * 4, mem, st);
    const int* ld = lens ? v->lens_dev : nullptr;
    long long c0 = launch_count();
    if (v->ctx.mode == F5B_MODE_BF16) v->run<act16>(st, B, T, ld);
    else v->run<float>(st, B, T, ld);
    set_last_call(launch_count() - c0);
    copy_out(wav, v->wav, (size_t)B * (T - 1) * 256 * 4, mem, st);
    F5B_CHECK_CUDA(cudaEventRecord(v->done, st));
    if (mem == F5B_MEM_HOST || sync_debug()) F5B_CHECK_CUDA(cudaStreamSynchronize(st));
  } catch (const f5b::Err& e) {
    return e.code;
  }
  return F5B_OK;
}

// Kernel-level entry (parity tests): the fused ISTFTHead tail on its own.  z [B * T, ldz] device fp32 (log-magnitude |
// phase, the output of head.out), lens [B] device int32 or NULL, wav [B, (T - 1) * 256] device.  Synchronous.
int f5b_op_istft(const float* z, int64_t ldz, const int32_t* lens, int32_t B, int32_t T, float* wav, int32_t device,
                 void* stream) {
  try {
    F5B_REQUIRE(z && wav && B > 0 && T > 1 && ldz >= V_OUT, "bad argument");
    open_device(device, F5B_MODE_FP32);
    stft_prepare(device);
    cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
    istft_run(st, device, z, ldz, B, T, lens, wav);
    F5B_CHECK_CUDA(cudaStreamSynchronize(st));
  } catch (const f5b::Err& e) {
    return e.code;
  }
  return F5B_OK;
}

int f5b_vocos_decode(f5b_vocos* v, const float* mel, int32_t B, int32_t T, float* wav, int32_t mem, void* stream) {
  return f5b_vocos_decode_varlen(v, mel, nullptr, B, T, wav, mem, stream);
}

}  // extern "C"
