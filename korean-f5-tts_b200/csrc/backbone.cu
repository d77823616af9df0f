// f5b_plan: the DiT / UNetT backbone, the text-embedding cache and the CFG + Euler sampling loop.
//
// Mirrors (by behaviour, not by structure) reference src/f5_tts/model/cfm.py:160-219 (ODE loop, CFG closure),
// backbones/dit.py:234-329, backbones/unett.py:192-280 and the modules they call (modules.py).  Everything that does
// not depend on the ODE state is hoisted out of the NFE loop: the time embedding and all AdaLN modulation vectors
// for every step are one skinny GEMM, the text embeddings are computed once per call, the cond/text columns of the
// input-projection operand are written once.  The whole call is captured into one CUDA graph per shape key.
#include <algorithm>
#include <cmath>
#include <cstdlib>
#include <cstring>
#include <mutex>

#include "plan.h"

namespace f5b {

DeviceCtx open_device(int device, int mode);

__global__ void conv_w_rearrange_f32(const float* __restrict__ w, float* __restrict__ out, int d, int cg, int taps) {
  // w [d, cg, taps] -> out [g][tap][ci][co_l]
  long long i = (long long)blockIdx.x * 256 + threadIdx.x;
  long long total = (long long)d * cg * taps;
  if (i >= total) return;
  int co_l = (int)(i % cg);
  long long r = i / cg;
  int ci = (int)(r % cg);
  r /= cg;
  int tap = (int)(r % taps);
  int g = (int)(r / taps);
  int co = g * cg + co_l;
  out[i] = w[((long long)co * cg + ci) * taps + tap];
}
__global__ void conv_w_rearrange_f16(const float* __restrict__ w, __half* __restrict__ out, int d, int cg, int taps) {
  // w [d, cg, taps] -> out [co][tap*cg + ci]
  long long i = (long long)blockIdx.x * 256 + threadIdx.x;
  long long total = (long long)d * cg * taps;
  if (i >= total) return;
  int ci = (int)(i % cg);
  long long r = i / cg;
  int tap = (int)(r % taps);
  int co = (int)(r / taps);
  out[i] = __float2half_rn(w[((long long)co * cg + ci) * taps + tap]);
}
__global__ void dw_transpose(const float* __restrict__ w, float* __restrict__ out, int d) {
  int i = blockIdx.x * 256 + threadIdx.x;
  if (i >= d * 7) return;
  int c = i / 7, k = i - c * 7;
  out[k * d + c] = w[i];
}
__global__ void dup_lens_kernel(const int* __restrict__ lens, int B, int batches, int add, int* __restrict__ out) {
  int i = blockIdx.x * 256 + threadIdx.x;
  if (i < batches) out[i] = lens[i % B] + add;
}

struct TextBlock {
  const float *dw_t, *dw_b, *ln_w, *ln_b, *b1, *gamma, *beta, *b2;
  const float* dw_w = nullptr;  // depthwise weight in the checkpoint's own [td, 1, 7] layout (backward pass)
  WMat pw1, pw2;
};
struct Layer {
  WMat qkv, wo, ff1, ff2, skip;
  const float *bqkv, *bo, *bff1, *bff2, *g1 = nullptr, *g3 = nullptr;
  bool has_skip = false;
};

struct Bufs {
  long long cap_rows = 0;  // capacities, each for one quantity (see ensure_ws): 2 B R, B N, B nt, table rows
  long long cap_bn = 0, cap_bnt = 0, cap_nT = 0;
  int capB = 0, capR = 0;
  float *y, *y2, *cond, *v_out, *x, *hproj, *hconst, *a_in, *te, *sin_emb, *h1, *st_emb, *ada, *rope_cs, *t_dev, *grn_scratch, *traj;
  long long traj_cap = 0;
  long long* text;
  int *lens_in, *lens2, *lens2p1;
  unsigned char* keep;
  void *c_in, *c_mid, *norm, *norm3, *qkv, *attn_o, *ffh, *tnorm, *th, *cat;
  std::vector<float*> skips;
};

struct GraphKey {
  int B = -1, N = 0, nt = 0, steps = 0, has_lens = 0, want_traj = 0;
  float cfg = 0;
  std::vector<float> t;
  bool operator==(const GraphKey& o) const {
    return B == o.B && N == o.N && nt == o.nt && steps == o.steps && has_lens == o.has_lens &&
           want_traj == o.want_traj && cfg == o.cfg && t == o.t;
  }
};

}  // namespace f5b

using namespace f5b;

struct f5b_plan {
  DeviceCtx ctx;
  f5b_arch a;
  Arena warena;
  Arena train_ws;  // workspace of the training step (backbone_train.inc), kept between steps
  std::unique_ptr<WeightStore> ws;
  std::unique_ptr<Arena> wk;  // workspace arena
  Bufs bf;
  int inner, R_extra, kin, kin_ld, td, pe_heads;
  bool is_dit;
  std::vector<Layer> layers;
  std::vector<TextBlock> tblocks;
  WMat in_proj, proj_out;
  const float *b_in, *b_out, *tm0_w, *tm0_b, *tm2_w, *tm2_b, *emb_table, *norm_out_g = nullptr;
  const float* conv_b[2];
  const float* conv_w_r[2] = {nullptr, nullptr};
  const float* conv_w_orig[2] = {nullptr, nullptr};  // [d, d/16, 31] as in the checkpoint (backward pass)
  const __half* conv_w16[2] = {nullptr, nullptr};  // fp16: see input_embed
  float *ada_w = nullptr, *ada_b = nullptr;
  int ada_W = 0;
  float *inv_freq_text, *inv_freq_rope, *time_freq;
  int* bad_id_dev = nullptr;   // set by text_lookup on a token id outside the embedding table
  int* bad_id_host = nullptr;  // pinned mirror, read once the call that wrote it has completed
  // Calls with device-resident buffers return as soon as their work is enqueued (like every PyTorch CUDA op): `done`
  // marks the end of the last call (the next call on ANY stream orders itself after it: the workspace is shared), a
  // bad token id of an asynchronous call is reported by the next call / f5b_plan_sync once `done` has completed.
  cudaEvent_t done = nullptr, fork = nullptr;
  bool pending = false;
  std::mutex mu;               // one call at a time per plan (the reference calls sample() from a thread pool)
  // forward-API text cache
  bool cache_valid = false;
  int cache_key[6] = {0, 0, 0, 0, 0, 0};
  // graph cache
  // instantiated sampling graphs, least recently used first; all of them point into the current arena
  struct GraphEntry { GraphKey key; cudaGraphExec_t exec; long long nodes; };
  std::vector<GraphEntry> gcache;
  static constexpr size_t kGraphCacheMax = 8;
  long long last_call_launches = 0;
  bool use_graph = true;
  cudaStream_t own_stream = nullptr;  // used when the caller passes the (uncapturable) legacy default stream
  float* save_x_last = nullptr;       // backward pass: eval() copies every block's input here ([depth, M, d])

  f5b_plan(const f5b_arch& arch, const f5b_tensor* w, int n, int mode, int device);
  ~f5b_plan();
  void build_weights(const f5b_tensor* w, int n, cudaStream_t st);  // copies + packs every weight into `warena`
  void refresh_weights(const f5b_tensor* w, int n);                 // the same again over the same buffers
  void ensure_ws(int B, int N, int nt, int steps, bool want_traj);
  template <typename T> void text_embed(cudaStream_t st, int B, int N, int nt, int batches, int text_uncond_from,
                                        int cond_uncond_from, const int* lens_text);
  void const_inputs(cudaStream_t st, int B, int batches, int N, int cond_uncond_from);
  template <typename T> void tables(cudaStream_t st, int n_t);
  template <typename T> void input_embed(cudaStream_t st, int batches, int N, const int* lens2);
  template <typename T> void eval(cudaStream_t st, int B, int N, int batches, int tbl_row, int tbl_bmod,
                                  const int* lens2, const int* lens2p1);
  template <typename T> void sample_body(cudaStream_t st, int B, int N, int nt, bool has_lens, const float* t_host,
                                         int steps, float cfg, bool want_traj);
  void drop_graph() {
    for (auto& e : gcache) cudaGraphExecDestroy(e.exec);
    gcache.clear();
  }
};

f5b_plan::~f5b_plan() {
  drop_graph();
  if (bad_id_host) cudaFreeHost(bad_id_host);
  if (done) cudaEventDestroy(done);
  if (fork) cudaEventDestroy(fork);
  if (own_stream) cudaStreamDestroy(own_stream);
}

void f5b_plan::build_weights(const f5b_tensor* w, int n, cudaStream_t st) {
  const int mode = ctx.mode;
  const int d = a.dim, ffd = a.dim * a.ff_mult;
  tblocks.clear();
  layers.clear();
  ws.reset(new WeightStore(&warena, w, n, st));
  auto W = [&](const std::string& nm, long long numel) { return ws->get(nm, numel); };

  tm0_w = W("time_embed.time_mlp.0.weight", (long long)d * 256);
  tm0_b = W("time_embed.time_mlp.0.bias", d);
  tm2_w = W("time_embed.time_mlp.2.weight", (long long)d * d);
  tm2_b = W("time_embed.time_mlp.2.bias", d);
  emb_table = W("text_embed.text_embed.weight", (long long)(a.text_num_embeds + 1) * td);
  for (int i = 0; i < a.conv_layers; ++i) {
    std::string p = "text_embed.text_blocks." + std::to_string(i);
    TextBlock tb;
    float* dwt = warena.alloc_n<float>((size_t)7 * td);
    dw_transpose<<<(td * 7 + 255) / 256, 256, 0, st>>>(W(p + ".dwconv.weight", (long long)td * 7), dwt, td);
    F5B_POST_LAUNCH();
    tb.dw_t = dwt;
    tb.dw_w = W(p + ".dwconv.weight", (long long)td * 7);
    tb.dw_b = W(p + ".dwconv.bias", td);
    tb.ln_w = W(p + ".norm.weight", td);
    tb.ln_b = W(p + ".norm.bias", td);
    tb.pw1 = make_wmat(ctx, &warena, st, W(p + ".pwconv1.weight", (long long)2 * td * td), 2 * td, td);
    tb.b1 = W(p + ".pwconv1.bias", 2 * td);
    tb.gamma = W(p + ".grn.gamma", 2 * td);
    tb.beta = W(p + ".grn.beta", 2 * td);
    tb.pw2 = make_wmat(ctx, &warena, st, W(p + ".pwconv2.weight", (long long)2 * td * td), td, 2 * td);
    tb.b2 = W(p + ".pwconv2.bias", td);
    tblocks.push_back(tb);
  }
  in_proj.f32 = W("input_embed.proj.weight", (long long)d * kin);  // contracted in fp32 in both modes (const_inputs)
  in_proj.N = d;
  in_proj.K = kin;
  in_proj.ld32 = kin;
  b_in = W("input_embed.proj.bias", d);
  const int cg = d / 16;
  for (int c = 0; c < 2; ++c) {
    std::string p = "input_embed.conv_pos_embed.conv1d." + std::to_string(c * 2);
    const float* cw = W(p + ".weight", (long long)d * cg * 31);
    conv_b[c] = W(p + ".bias", d);
    conv_w_orig[c] = cw;
    long long total = (long long)d * cg * 31;
    if (mode == F5B_MODE_BF16) {
      __half* o = warena.alloc_n<__half>((size_t)total);
      conv_w_rearrange_f16<<<(unsigned)((total + 255) / 256), 256, 0, st>>>(cw, o, d, cg, 31);
      conv_w16[c] = o;
    } else {
      float* o = warena.alloc_n<float>((size_t)total);
      conv_w_rearrange_f32<<<(unsigned)((total + 255) / 256), 256, 0, st>>>(cw, o, d, cg, 31);
      conv_w_r[c] = o;
    }
    F5B_POST_LAUNCH();
  }
  std::vector<std::pair<const float*, long long>> ada_ws, ada_bs;
  for (int i = 0; i < a.depth; ++i) {
    Layer L;
    std::string pa, pf;
    if (is_dit) {
      std::string p = "transformer_blocks." + std::to_string(i);
      pa = p + ".attn";
      pf = p + ".ff";
      ada_ws.push_back({W(p + ".attn_norm.linear.weight", (long long)6 * d * d), (long long)6 * d * d});
      ada_bs.push_back({W(p + ".attn_norm.linear.bias", 6 * d), 6 * d});
    } else {
      std::string p = "layers." + std::to_string(i);
      pa = p + ".2";
      pf = p + ".4";
      L.g1 = W(p + ".1.g", d);
      L.g3 = W(p + ".3.g", d);
      if (i >= a.depth / 2) {
        L.has_skip = true;
        L.skip = make_wmat(ctx, &warena, st, W(p + ".0.weight", (long long)2 * d * d), d, 2 * d);
      }
    }
    const long long qn = (long long)inner * d;
    float* wqkv = concat_rows(&warena, st, {{W(pa + ".to_q.weight", qn), qn}, {W(pa + ".to_k.weight", qn), qn},
                                            {W(pa + ".to_v.weight", qn), qn}});
    L.bqkv = concat_rows(&warena, st, {{W(pa + ".to_q.bias", inner), inner}, {W(pa + ".to_k.bias", inner), inner},
                                       {W(pa + ".to_v.bias", inner), inner}});
    L.qkv = make_wmat(ctx, &warena, st, wqkv, 3 * inner, d);
    L.wo = make_wmat(ctx, &warena, st, W(pa + ".to_out.0.weight", qn), d, inner);
    L.bo = W(pa + ".to_out.0.bias", d);
    L.ff1 = make_wmat(ctx, &warena, st, W(pf + ".ff.0.0.weight", (long long)ffd * d), ffd, d);
    L.bff1 = W(pf + ".ff.0.0.bias", ffd);
    L.ff2 = make_wmat(ctx, &warena, st, W(pf + ".ff.2.weight", (long long)ffd * d), d, ffd);
    L.bff2 = W(pf + ".ff.2.bias", d);
    layers.push_back(L);
  }
  if (is_dit) {
    ada_ws.push_back({W("norm_out.linear.weight", (long long)2 * d * d), (long long)2 * d * d});
    ada_bs.push_back({W("norm_out.linear.bias", 2 * d), 2 * d});
    ada_w = concat_rows(&warena, st, ada_ws);
    ada_b = concat_rows(&warena, st, ada_bs);
    ada_W = a.depth * 6 * d + 2 * d;
  } else {
    norm_out_g = W("norm_out.g", d);
  }
  // the velocity goes straight into the ODE state: proj_out is contracted with ~16-bit operands in bf16 mode
  proj_out = make_wmat_split3(ctx, &warena, st, W("proj_out.weight", (long long)a.mel_dim * d), a.mel_dim, d);
  b_out = W("proj_out.bias", a.mel_dim);
}

// After an optimiser step (f5b_plan_refresh): the SAME sequence of copies / packing kernels over the SAME device
// buffers (Arena replay), so the weights change under the plan without a single cudaMalloc / cudaFree, the workspace
// and the captured graphs (which hold pointers into these buffers) stay valid.
void f5b_plan::refresh_weights(const f5b_tensor* w, int n) {
  F5B_CHECK_CUDA(cudaSetDevice(ctx.device));
  std::lock_guard<std::mutex> lock(mu);
  F5B_CHECK_CUDA(cudaStreamSynchronize(own_stream));  // nothing of an earlier call may still read the old values
  const size_t before = warena.blocks();
  warena.begin_replay();
  try {
    build_weights(w, n, 0);
  } catch (...) {
    warena.end_replay();
    throw;
  }
  warena.end_replay();
  F5B_REQUIRE(warena.blocks() == before, "the refreshed state dict does not have the shape of the plan's");
  cache_valid = false;
  F5B_CHECK_CUDA(cudaStreamSynchronize(0));
}

f5b_plan::f5b_plan(const f5b_arch& arch, const f5b_tensor* w, int n, int mode, int device) : a(arch) {
  F5B_REQUIRE(mode == F5B_MODE_FP32 || mode == F5B_MODE_BF16, "unknown precision mode");
  F5B_REQUIRE(a.backbone == F5B_BACKBONE_DIT || a.backbone == F5B_BACKBONE_UNETT, "unknown backbone");
  F5B_REQUIRE(a.dim_head == 64, "only dim_head 64 is built (all shipped configs)");
  F5B_REQUIRE(a.dim % 16 == 0 && a.mel_dim % 4 == 0 && a.text_dim % 4 == 0, "unsupported widths");
  ctx = open_device(device, mode);
  use_graph = std::getenv("F5B_NO_GRAPH") == nullptr;
  F5B_CHECK_CUDA(cudaStreamCreateWithFlags(&own_stream, cudaStreamNonBlocking));
  is_dit = a.backbone == F5B_BACKBONE_DIT;
  inner = a.heads * 64;
  R_extra = is_dit ? 0 : 1;
  td = a.text_dim;
  kin = 2 * a.mel_dim + td;
  kin_ld = (kin + 7) / 8 * 8;
  pe_heads = a.pe_attn_head > 0 ? a.pe_attn_head : a.heads;
  const int d = a.dim, ffd = a.dim * a.ff_mult;
  if (mode == F5B_MODE_BF16) F5B_REQUIRE((d / 16) % 64 == 0, "bf16 mode needs dim/16 channels per conv group to be a multiple of 64");

  cudaStream_t st = 0;
  build_weights(w, n, st);
  // frequency tables, computed the way the reference does (fp32 pow / exp of fp32 arguments)
  std::vector<float> ft(td / 2), fr(32), tf(128);
  for (int j = 0; j < td / 2; ++j) ft[j] = 1.0f / powf(10000.0f, (float)(2 * j) / (float)td);          // modules.py:202-213
  for (int j = 0; j < 32; ++j) fr[j] = 1.0f / powf(10000.0f, (float)(2 * j) / 64.0f);                   // x_transformers RoPE
  const float kk = (float)(-(std::log(10000.0) / 127.0));
  for (int j = 0; j < 128; ++j) tf[j] = expf((float)j * kk);                                            // modules.py:158-160
  inv_freq_text = warena.alloc_n<float>(ft.size() ? ft.size() : 1);
  inv_freq_rope = warena.alloc_n<float>(32);
  time_freq = warena.alloc_n<float>(128);
  bad_id_dev = warena.alloc_n<int>(1);
  F5B_CHECK_CUDA(cudaMemset(bad_id_dev, 0, sizeof(int)));
  F5B_CHECK_CUDA(cudaHostAlloc(&bad_id_host, sizeof(int), cudaHostAllocDefault));
  *bad_id_host = 0;
  F5B_CHECK_CUDA(cudaEventCreateWithFlags(&done, cudaEventDisableTiming));
  F5B_CHECK_CUDA(cudaEventCreateWithFlags(&fork, cudaEventDisableTiming));
  if (!ft.empty()) F5B_CHECK_CUDA(cudaMemcpy(inv_freq_text, ft.data(), ft.size() * 4, cudaMemcpyHostToDevice));
  F5B_CHECK_CUDA(cudaMemcpy(inv_freq_rope, fr.data(), 32 * 4, cudaMemcpyHostToDevice));
  F5B_CHECK_CUDA(cudaMemcpy(time_freq, tf.data(), 128 * 4, cudaMemcpyHostToDevice));
  F5B_CHECK_CUDA(cudaStreamSynchronize(st));
}

// ---------------------------------------------------------------------------------------------
// workspace
// ---------------------------------------------------------------------------------------------
void f5b_plan::ensure_ws(int B, int N, int nt, int steps, bool want_traj) {
  // Capacities are tracked per QUANTITY (rows, B * N, B * nt, table rows, ...) and only ever grow: a job whose
  // micro-batches trade batch size against length (C3: B = 8 x N = 1874 ... B = 57 x N = 285, always ~16 k rows) used
  // to re-allocate the whole workspace -- device-wide synchronisation, ~45 cudaMalloc / cudaFree of up to hundreds of MB
  // -- at EVERY call, because one request-shaped key (B, N, ...) never covered the next one.
  const int R = N + R_extra;
  const long long rows = 2LL * B * R, bn = (long long)B * N, bnt = (long long)B * (nt > 0 ? nt : 1);
  const long long nT = steps > 2 * B ? steps : 2 * B;  // table rows: steps (sample) or B (forward)
  const long long traj_need = want_traj ? (long long)(steps + 1) * B * N * a.mel_dim : 0;
  if (wk && rows <= bf.cap_rows && bn <= bf.cap_bn && bnt <= bf.cap_bnt && B <= bf.capB && nT <= bf.cap_nT &&
      R <= bf.capR && traj_need <= bf.traj_cap)
    return;
  drop_graph();
  cache_valid = false;
  F5B_CHECK_CUDA(cudaDeviceSynchronize());
  Bufs nb;
  auto grow = [](long long need, long long have) {  // 1/8 headroom when growing: slightly longer requests fit
    return need <= have ? have : need + need / 8;
  };
  nb.cap_rows = grow(rows, wk ? bf.cap_rows : 0);
  nb.cap_bn = grow(bn, wk ? bf.cap_bn : 0);
  nb.cap_bnt = grow(bnt, wk ? bf.cap_bnt : 0);
  nb.capB = (int)grow(B, wk ? bf.capB : 0);
  nb.cap_nT = grow(nT, wk ? bf.cap_nT : 0);
  nb.capR = (int)grow(R, wk ? bf.capR : 0);
  nb.traj_cap = traj_need <= (wk ? bf.traj_cap : 0) ? bf.traj_cap : traj_need;
  // The old arena goes first (peak memory stays at one workspace); until every allocation below has succeeded the
  // plan holds NO workspace and zero capacities, so a failed growth (OOM on a large request) cannot leave stale
  // pointers behind a passing capacity check: the next call simply allocates again.
  wk.reset();
  this->bf = Bufs();
  std::unique_ptr<Arena> fresh(new Arena());
  Arena& A = *fresh;
  Bufs& bf = nb;  // (shadows the member on purpose: everything below fills the local copy)
  const int d = a.dim, mel = a.mel_dim, ffd = a.dim * a.ff_mult;
  const size_t es = ctx.mode == F5B_MODE_BF16 ? 2 : 4;
  const size_t crows = (size_t)bf.cap_rows, cbn = (size_t)bf.cap_bn, cT = (size_t)bf.cap_nT, cB = (size_t)bf.capB;
  bf.y = A.alloc_n<float>(cbn * mel);
  bf.y2 = A.alloc_n<float>(cbn * mel);
  bf.cond = A.alloc_n<float>(cbn * mel);
  bf.v_out = A.alloc_n<float>(crows * mel);  // UNetT keeps the time-token row; consumers skip it
  bf.x = A.alloc_n<float>(crows * d);
  bf.hproj = A.alloc_n<float>(crows * d);
  bf.te = A.alloc_n<float>(2 * cbn * td);
  bf.sin_emb = A.alloc_n<float>(cT * 256);
  bf.h1 = A.alloc_n<float>(cT * d);
  bf.st_emb = A.alloc_n<float>(cT * d);
  bf.ada = A.alloc_n<float>(cT * (ada_W > 0 ? ada_W : 1));
  bf.rope_cs = A.alloc_n<float>((size_t)bf.capR * 64);
  bf.t_dev = A.alloc_n<float>(cT + 1);
  bf.grn_scratch = A.alloc_n<float>(2 * cB * 2 * 2 * td);
  bf.traj = bf.traj_cap ? A.alloc_n<float>((size_t)bf.traj_cap) : nullptr;
  bf.text = A.alloc_n<long long>((size_t)bf.cap_bnt);
  bf.lens_in = A.alloc_n<int>(cB);
  bf.lens2 = A.alloc_n<int>(2 * cB);
  bf.lens2p1 = A.alloc_n<int>(2 * cB);
  bf.keep = A.alloc_n<unsigned char>(2 * cbn);
  bf.hconst = A.alloc_n<float>(2 * cbn * d);
  bf.a_in = A.alloc_n<float>(2 * cbn * kin_ld);
  F5B_CHECK_CUDA(cudaMemset(bf.a_in, 0, 2 * cbn * kin_ld * 4));
  bf.c_in = A.alloc(crows * d * es);
  bf.c_mid = A.alloc(crows * d * es);
  bf.norm = A.alloc(crows * d * es);
  bf.norm3 = ctx.mode == F5B_MODE_BF16 ? A.alloc(crows * 3 * d * 2) : nullptr;
  bf.qkv = A.alloc(crows * 3 * inner * es);
  bf.attn_o = A.alloc(crows * inner * es);
  bf.ffh = A.alloc(crows * ffd * es);
  bf.tnorm = A.alloc(2 * cbn * td * es);
  bf.th = A.alloc(2 * cbn * 2 * td * es);
  bf.skips.clear();
  bf.cat = nullptr;
  if (!is_dit) {
    bf.cat = A.alloc(crows * 2 * d * es);
    for (int i = 0; i < a.depth / 2; ++i) bf.skips.push_back(A.alloc_n<float>(crows * d));
  }
  this->bf = nb;  // commit: every allocation succeeded
  wk = std::move(fresh);
}

// ---------------------------------------------------------------------------------------------
// text embedding (once per sample() call) + assembly of the constant columns of A_in
// ---------------------------------------------------------------------------------------------
template <typename T>
void f5b_plan::text_embed(cudaStream_t st, int B, int N, int nt, int batches, int text_uncond_from,
                          int cond_uncond_from, const int* lens_text) {
  const bool extra = a.conv_layers > 0;
  text_lookup(st, bf.text, B, nt, text_uncond_from, batches, N, td, lens_text, emb_table,
              extra ? inv_freq_text : nullptr, a.text_mask_padding, bf.te, bf.keep, a.text_num_embeds + 1, bad_id_dev);
  const long long M = (long long)batches * N;
  const bool need_mask = extra && (a.text_mask_padding || lens_text != nullptr);
  for (size_t i = 0; i < tblocks.size(); ++i) {
    const TextBlock& tb = tblocks[i];
    dwconv7_ln<T>(st, bf.te, (T*)bf.tnorm, batches, N, td, lens_text, tb.dw_t, tb.dw_b, tb.ln_w, tb.ln_b, 1e-6f);
    EpiParams e1;
    e1.act = ACT_GELU_ERF;
    e1.bias = tb.b1;
    e1.out = bf.th;
    e1.out_type = sizeof(T) == 2 ? OUT_F16 : OUT_F32;
    e1.ldo = 2 * td;
    e1.n_valid = 2 * td;
    run_gemm(ctx, st, bf.tnorm, td, M, tb.pw1, e1);
    grn_inplace<T>(st, (T*)bf.th, batches, N, 2 * td, lens_text, tb.gamma, tb.beta, bf.grn_scratch);
    EpiParams e2;
    e2.kind = EPI_RESID;
    e2.bias = tb.b2;
    e2.resid = bf.te;
    e2.ldr = td;
    e2.n_valid = td;
    run_gemm(ctx, st, bf.th, 2 * td, M, tb.pw2, e2);
    if (need_mask) mask_rows(st, bf.te, M, td, bf.keep);
  }
  const_inputs(st, B, batches, N, cond_uncond_from);
}

// The input projection (dit.py:124,138 / unett.py:92,99) is linear in cat(x, cond, text): the cond / text columns do
// not change during the ODE, so their contribution (+ bias) is contracted once per call, in fp32, into bf.hconst;
// every evaluation only adds the K = mel_dim x-columns (fp32 as well: the ODE state never passes through bf16).
void f5b_plan::const_inputs(cudaStream_t st, int B, int batches, int N, int cond_uncond_from) {
  const int mel = a.mel_dim, d = a.dim;
  build_a_in(st, bf.a_in, kin_ld, B, batches, N, mel, td, bf.cond, bf.te, cond_uncond_from);
  EpiParams e;
  e.bias = b_in;
  e.out = bf.hconst;
  e.ldo = d;
  e.n_valid = d;
  simt_gemm(st, bf.a_in + mel, kin_ld, (long long)batches * N, in_proj.f32 + mel, kin, d, kin - mel, e);
}

// ---------------------------------------------------------------------------------------------
// per-call tables: time embedding and every AdaLN modulation vector for n_t time values (bf.t_dev)
// ---------------------------------------------------------------------------------------------
template <typename T>
void f5b_plan::tables(cudaStream_t st, int n_t) {
  const int d = a.dim;
  sinus_time(st, bf.t_dev, n_t, time_freq, bf.sin_emb);
  EpiParams e;
  e.act = ACT_SILU;
  e.bias = tm0_b;
  e.out = bf.h1;
  e.ldo = d;
  e.n_valid = d;
  simt_gemm(st, bf.sin_emb, 256, n_t, tm0_w, 256, d, 256, e);
  EpiParams e2;
  e2.act = is_dit ? ACT_SILU : ACT_NONE;  // DiT consumes t only through SiLU (modules.py:316,337); UNetT uses t raw
  e2.bias = tm2_b;
  e2.out = bf.st_emb;
  e2.ldo = d;
  e2.n_valid = d;
  simt_gemm(st, bf.h1, d, n_t, tm2_w, d, d, d, e2);
  if (is_dit) {
    EpiParams e3;
    e3.bias = ada_b;
    e3.out = bf.ada;
    e3.ldo = ada_W;
    e3.n_valid = ada_W;
    simt_gemm(st, bf.st_emb, d, n_t, ada_w, d, ada_W, d, e3);
  }
}

// ---------------------------------------------------------------------------------------------
// InputEmbedding: proj + ConvPositionEmbedding + residual  ->  bf.x (DiT) / bf.hproj-as-h (UNetT)
// ---------------------------------------------------------------------------------------------
template <typename T>
void f5b_plan::input_embed(cudaStream_t st, int batches, int N, const int* lens_conv) {
  const int d = a.dim;
  const long long M = (long long)batches * N;
  const bool is16 = sizeof(T) == 2;
  // bf.hproj = hconst + y W_x^T was written by euler_xproj (fused with the previous Euler update)
  // bf16 mode feeds the position-embedding conv IEEE fp16 operands (kind::f16 runs both formats at the same rate):
  // h carries the large, slowly varying cond/text offsets and loses too much in 8 mantissa bits (precision)
  float* xdst = is_dit ? bf.x : bf.skips[0];
  if (is16) cast_mask<__half>(st, bf.hproj, (__half*)bf.c_in, batches, N, d, lens_conv, xdst);  // xdst = h
  else cast_mask<float>(st, bf.hproj, (float*)bf.c_in, batches, N, d, lens_conv);
  for (int c = 0; c < 2; ++c) {
    EpiParams ec;
    ec.act = ACT_MISH;
    ec.bias = conv_b[c];
    ec.n_valid = d;
    ec.rows_per_batch = N;
    ec.row_len = lens_conv;
    const void* src = c == 0 ? bf.c_in : bf.c_mid;
    if (c == 0) {
      ec.out = bf.c_mid;
      ec.out_type = is16 ? OUT_F16 : OUT_F32;
      ec.ldo = d;
    } else if (is16) {
      // x = h + mish(conv): xdst already holds h, the tensor-core epilogue adds the conv branch (TMA reduce-add)
      ec.kind = EPI_RESID;
      ec.resid = xdst;
      ec.ldr = d;
    } else {
      ec.out = xdst;
      ec.out_type = OUT_F32;
      ec.ldo = d;
      ec.addend = bf.hproj;
      ec.ld_add = d;
    }
    if (is16) {
      TcConv cv{31, 15, d / 16, d, 1};
      tc_gemm(st, ctx.sm_count, (const bf16*)src, d, N, batches, (const bf16*)conv_w16[c], (long long)31 * (d / 16), d,
              31 * (d / 16), ec, &cv, 0);
    } else {
      simt_group_conv(st, (const float*)src, batches, N, d, conv_w_r[c], 31, 15, 16, ec);
    }
  }
}

// ---------------------------------------------------------------------------------------------
// one backbone evaluation: bf.a_in -> bf.v_out [batches*R, mel]  (R = N, or N + 1 with UNetT's leading time token)
// ---------------------------------------------------------------------------------------------
template <typename T>
void f5b_plan::eval(cudaStream_t st, int B, int N, int batches, int tbl_row, int tbl_bmod, const int* lens2,
                    const int* lens2p1) {
  const int d = a.dim, ffd = a.dim * a.ff_mult, mel = a.mel_dim;
  const int R = N + R_extra;
  const long long M = (long long)batches * R;
  const bool is16 = sizeof(T) == 2;
  const long long tb_stride = tbl_bmod > 1 ? ada_W : 0;
  const int* lensR = is_dit ? lens2 : lens2p1;
  const int* kv_len = a.attn_mask_enabled ? lensR : nullptr;

  input_embed<T>(st, batches, N, is_dit ? lens2 : nullptr);
  if (!is_dit) unett_prepend(st, bf.st_emb + (long long)tbl_row * d, bf.skips[0], bf.x, batches, N, d, tbl_bmod);

  int skip_top = 0;
  std::vector<float*> stack;
  for (int l = 0; l < a.depth; ++l) {
    const Layer& L = layers[l];
    if (save_x_last != nullptr) {
      F5B_CHECK_CUDA(cudaMemcpyAsync(save_x_last + (size_t)l * M * d, bf.x, (size_t)M * d * 4, cudaMemcpyDeviceToDevice, st));
      count_launch(1);
    }
    const float* ada = is_dit ? bf.ada + (long long)tbl_row * ada_W + (long long)l * 6 * d : nullptr;
    if (!is_dit) {
      if (l < a.depth / 2) {
        // skips[0] doubles as the input-embed scratch above; it is consumed before layer 0 overwrites it
        F5B_CHECK_CUDA(cudaMemcpyAsync(bf.skips[skip_top], bf.x, (size_t)M * d * 4, cudaMemcpyDeviceToDevice, st));
        count_launch(1);
        stack.push_back(bf.skips[skip_top++]);
      } else {
        float* sk = stack.back();
        stack.pop_back();
        concat2<T>(st, bf.x, sk, (T*)bf.cat, M, d);
        EpiParams es;
        es.out = bf.x;
        es.ldo = d;
        es.n_valid = d;
        run_gemm(ctx, st, bf.cat, 2 * d, M, L.skip, es);
      }
    }
    // attention branch
    if (is_dit)
      ln_mod<T>(st, bf.x, (T*)bf.norm, M, d, ada + d, ada, 1.0f, tb_stride, R, tbl_bmod, 1e-6f);
    else
      xrms_norm<T>(st, bf.x, (T*)bf.norm, M, d, L.g1);
    EpiParams eq;
    eq.bias = L.bqkv;
    eq.out = bf.qkv;
    eq.out_type = is16 ? OUT_BF16 : OUT_F32;  // q, k, v feed the attention kernel, which contracts bf16
    eq.ldo = 3 * inner;
    eq.n_valid = 3 * inner;
    eq.rows_per_batch = R;
    eq.rope_cols = 2 * inner;
    eq.rope_period = inner;
    eq.rope_heads = pe_heads;
    eq.rope_cs = bf.rope_cs;
    run_gemm(ctx, st, bf.norm, d, M, L.qkv, eq);
    if (is16)
      tc_attention(st, (const bf16*)bf.qkv, (bf16*)bf.attn_o, batches, R, a.heads, kv_len, 0.125f, 1);
    else
      simt_attention(st, (const float*)bf.qkv, (float*)bf.attn_o, batches, R, a.heads, kv_len, 0.125f);
    EpiParams eo;
    eo.kind = EPI_RESID;
    eo.bias = L.bo;
    eo.resid = bf.x;
    eo.ldr = d;
    eo.n_valid = d;
    eo.rows_per_batch = R;
    eo.row_len = lensR;  // masked_fill(~mask, 0) after to_out (modules.py:540-542)
    if (is_dit) {
      eo.gate = ada + 2 * d;
      eo.gate_bstride = tb_stride;
      eo.gate_bmod = tbl_bmod;
    }
    run_gemm(ctx, st, bf.attn_o, inner, M, L.wo, eo);
    // feed-forward branch
    if (is_dit)
      ln_mod<T>(st, bf.x, (T*)bf.norm, M, d, ada + 4 * d, ada + 3 * d, 1.0f, tb_stride, R, tbl_bmod, 1e-6f);
    else
      xrms_norm<T>(st, bf.x, (T*)bf.norm, M, d, L.g3);
    EpiParams e1;
    e1.act = ACT_GELU_TANH;
    e1.bias = L.bff1;
    e1.out = bf.ffh;
    e1.out_type = is16 ? OUT_F16 : OUT_F32;
    e1.ldo = ffd;
    e1.n_valid = ffd;
    run_gemm(ctx, st, bf.norm, d, M, L.ff1, e1);
    EpiParams e2;
    e2.kind = EPI_RESID;
    e2.bias = L.bff2;
    e2.resid = bf.x;
    e2.ldr = d;
    e2.n_valid = d;
    e2.rows_per_batch = R;
    if (is_dit) {
      e2.gate = ada + 5 * d;
      e2.gate_bstride = tb_stride;
      e2.gate_bmod = tbl_bmod;
    }
    run_gemm(ctx, st, bf.ffh, ffd, M, L.ff2, e2);
  }
  EpiParams ep;
  ep.bias = b_out;
  ep.out = bf.v_out;
  ep.ldo = mel;
  ep.n_valid = mel;
  ep.rows_per_batch = R;
  if (is_dit) {
    const float* fin = bf.ada + (long long)tbl_row * ada_W + (long long)a.depth * 6 * d;  // (scale, shift)
    if (is16) ln_mod<bf16x3>(st, bf.x, (bf16x3*)bf.norm3, M, d, fin, fin + d, 1.0f, tb_stride, R, tbl_bmod, 1e-6f);
    else ln_mod<T>(st, bf.x, (T*)bf.norm, M, d, fin, fin + d, 1.0f, tb_stride, R, tbl_bmod, 1e-6f);
  } else {
    if (is16) xrms_norm<bf16x3>(st, bf.x, (bf16x3*)bf.norm3, M, d, norm_out_g);
    else xrms_norm<T>(st, bf.x, (T*)bf.norm, M, d, norm_out_g);
  }
  if (is16) run_gemm(ctx, st, bf.norm3, 3 * d, M, proj_out, ep, 64);
  else run_gemm(ctx, st, bf.norm, d, M, proj_out, ep, 64);
}

#include "backbone_api.inc"
#include "backbone_train.inc"
