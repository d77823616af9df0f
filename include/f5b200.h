/* libf5b200 -- C ABI of the B200-native CFM.sample + Vocos hot path.
 *
 * The reference (KwangsungKim138/Korean-F5-TTS, a fork of SWivid/F5-TTS 1.1.15) has no FFI/plugin interface: its
 * "operator API" for this path is Python.  Each entry point below replaces the body of one reference
 * Python callable; the Python mirror in korean-f5-tts_b200/f5b200/ keeps the reference signatures and binds these
 * symbols with ctypes (INTEGRATION.md shows the stub a reference maintainer would add).
 *
 * Conventions
 *   - plain pointers and sizes only; tensors are dense row-major fp32 unless stated; `text` is int64 with -1 padding.
 *   - `mem` says where the data pointers live: F5B_MEM_HOST (the library stages H2D/D2H itself, pinned staging is the
 *     caller's business) or F5B_MEM_DEVICE (pointers on the plan's CUDA device).  `stream` is a cudaStream_t (may be 0).
 *   - every function returns 0 or a negative F5B_ERR_* code; f5b_last_error() gives the thread-local message.
 *   - there is NO CPU fallback: without an sm_100 device every compute entry returns F5B_ERR_NO_DEVICE.
 */
#ifndef F5B200_H_
#define F5B200_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define F5B_OK 0
#define F5B_ERR_INVALID (-1)
#define F5B_ERR_CUDA (-2)
#define F5B_ERR_NO_DEVICE (-3)
#define F5B_ERR_MISSING (-4)
#define F5B_ERR_INDEX (-5) /* a text id outside [-1, text_num_embeds): nn.Embedding raises IndexError (dit.py:42,94) */

#define F5B_MEM_HOST 0
#define F5B_MEM_DEVICE 1

#define F5B_MODE_FP32 0 /* CUDA-core fp32 contractions: parity mode (mel/wave max-abs 1e-3 vs the fp32 reference) */
#define F5B_MODE_BF16 1 /* tcgen05 bf16 contractions, fp32 accumulate / residual / ODE state: performance mode */

#define F5B_BACKBONE_DIT 0   /* reference src/f5_tts/model/backbones/dit.py:146-329 */
#define F5B_BACKBONE_UNETT 1 /* reference src/f5_tts/model/backbones/unett.py:107-280 */

/* Constructor arguments of DiT (dit.py:147-168) / UNetT (unett.py:108-127) that change the arithmetic. */
typedef struct f5b_arch {
  int32_t backbone;
  int32_t dim, depth, heads, dim_head, ff_mult, mel_dim;
  int32_t text_num_embeds; /* as passed to the constructor; the embedding table has text_num_embeds + 1 rows */
  int32_t text_dim, conv_layers;
  int32_t text_mask_padding;
  int32_t pe_attn_head; /* number of leading heads that get RoPE; <= 0 means all (None in the reference) */
  int32_t attn_mask_enabled;
} f5b_arch;

/* One state-dict entry (names exactly as the reference registers them, without the "transformer." prefix). */
typedef struct f5b_tensor {
  const char* name;
  const float* data; /* fp32, contiguous */
  int64_t numel;
  int32_t on_device; /* 0: host pointer, 1: device pointer */
} f5b_tensor;

typedef struct f5b_plan f5b_plan;   /* backbone weights + workspace + CUDA graphs (one per shape key) */
typedef struct f5b_vocos f5b_vocos; /* Vocos weights + workspace */

const char* f5b_last_error(void);
int f5b_version(void);
/* 0 if an sm_100 device is usable, F5B_ERR_NO_DEVICE otherwise (message says why). */
int f5b_device_check(int device);

/* Replaces: DiT.__init__/UNetT.__init__ + load_state_dict placement on device (infer/utils_infer.py:292-345). */
int f5b_plan_create(const f5b_arch* arch, const f5b_tensor* weights, int32_t n_weights, int32_t mode, int32_t device,
                    f5b_plan** out);
void f5b_plan_destroy(f5b_plan* plan);
/* After the optimiser has changed the weights (model/trainer.py:388 optimizer.step(); the reference's modules simply
 * read the updated nn.Parameters on the next forward): copy + re-pack the same state dict (same names and sizes as at
 * creation) into the plan's existing device buffers.  No allocation; workspace and graphs stay valid.  Synchronous:
 * waits for `stream` (the stream the update ran on) first and returns when the new weights are in place. */
int f5b_plan_refresh(f5b_plan* plan, const f5b_tensor* weights, int32_t n_weights, void* stream);
/* Grow the plan's workspace NOW for a batch of B utterances of N frames, nt text tokens, `steps` ODE steps (the
 * capacities are kept per quantity and never shrink).  A caller that knows the shapes of a whole job -- the chunks of
 * infer_batch_process (infer/utils_infer.py:500-770), the micro-batches of an evaluation run
 * (eval/eval_infer_batch.py) -- reserves them up front, so no call inside the job re-allocates (a re-allocation
 * synchronises the device and drops the cached graphs).  f5b_vocos_reserve does the same for Vocos (B x T frames). */
int f5b_plan_reserve(f5b_plan* plan, int32_t B, int32_t N, int32_t nt, int32_t steps);

/* Replaces the ODE part of CFM.sample (model/cfm.py:160-219): text-embedding cache, EPSS/sway grid consumption,
 * CFG-doubled backbone evaluations, Euler updates.  Inputs are what cfm.py:104-216 builds on the host:
 *   y0        [B, N, mel]   seeded noise (cfm.py:196-201)
 *   step_cond [B, N, mel]   where(cond_mask, cond, 0) (cfm.py:151-153)
 *   text      [B, nt] int64 token ids, -1 padded (model/utils.py:99-106)
 *   durations [B] int32 or NULL.  NULL <=> the reference's mask=None (batch 1, cfm.py:155-158)
 *   t_grid    [steps+1] HOST floats, already sway-warped (cfm.py:211-216)
 * Outputs: y_out [B, N, mel] = trajectory[-1]; traj [steps+1, B, N, mel] or NULL. */
int f5b_sample(f5b_plan* plan, const float* y0, const float* step_cond, const int64_t* text, const int32_t* durations,
               int32_t B, int32_t N, int32_t nt, const float* t_grid, int32_t steps, float cfg_strength, float* y_out,
               float* traj, int32_t mem, void* stream);

/* Replaces DiT.forward / UNetT.forward (dit.py:278-329, unett.py:217-280): one function evaluation.
 *   x, cond [B, N, mel]; time [B] HOST floats; durations as above; out [(cfg_infer ? 2B : B), N, mel].
 *   cache != 0 keeps the text embeddings until f5b_clear_cache (dit.py:259-276). */
int f5b_forward(f5b_plan* plan, const float* x, const float* cond, const int64_t* text, const float* time,
                const int32_t* durations, int32_t B, int32_t N, int32_t nt, int32_t drop_audio_cond, int32_t drop_text,
                int32_t cfg_infer, int32_t cache, float* out, int32_t mem, void* stream);
int f5b_clear_cache(f5b_plan* plan);
/* Calls whose buffers are on the device (mem = F5B_MEM_DEVICE) return once their work is enqueued on `stream`, like
 * any PyTorch CUDA operator; with host buffers they are synchronous.  f5b_plan_sync waits for the plan's last call and
 * returns its deferred status: F5B_ERR_INDEX when an asynchronous call met a text id outside the embedding table
 * (otherwise that error is returned by the next call on the plan, once the offending one has completed). */
int f5b_plan_sync(f5b_plan* plan);

/* Replaces the forward half of CFM.forward (model/cfm.py:231-302), i.e. the flow-matching loss of one training batch
 * after the random draws (which stay host-side Python like the reference's): phi = (1 - t) x0 + t x1, cond = x1 outside
 * the random span, one backbone evaluation with the drop flags, MSE against x1 - x0 over the span frames.
 *   x1, x0 [B, N, mel]; time [B] HOST floats; lens [B] int32 or NULL; span [B, N] uint8 = rand_span_mask & lens mask.
 * Outputs: loss [1]; cond, pred [B, N, mel] (either may be NULL).  Gradients: f5b_cfm_loss_tail_grads below. */
int f5b_cfm_loss(f5b_plan* plan, const float* x1, const float* x0, const int64_t* text, const float* time,
                 const int32_t* lens, const uint8_t* span, int32_t B, int32_t N, int32_t nt, int32_t drop_audio_cond,
                 int32_t drop_text, float* loss, float* cond, float* pred, int32_t mem, void* stream);

/* The backward pass of the training step (model/trainer.py:384 `accelerator.backward(loss)` over CFM.forward,
 * cfm.py:231-302; DiT backbone): same inputs as f5b_cfm_loss; the loss is differentiated through proj_out, the final
 * AdaLN, EVERY DiTBlock (dit.py:314-327, modules.py:683-697; each block recomputed from its saved input), the timestep
 * MLP, InputEmbedding and TextEmbedding.  Each gradient slot names a parameter by its state-dict name (no
 * "transformer." prefix), or "d_block_input" / "d_input_embed" [B, N, dim] (gradient of the residual stream entering the
 * last block / block 0), or "pred" [B, N, mel]; other names return F5B_ERR_MISSING.  fp32 mode: CUDA-core arithmetic
 * (parity tier); 16-bit mode: GEMMs, grouped convolutions and attention on tcgen05 with fp32 accumulation.  Gradients
 * are written as soon as their layer is done (last block first).  `ready_event`, when not NULL, is a cudaEvent_t that
 * the library records on its stream right after the slot has been written: slots of one all-reduce bucket share an
 * event, and a communication stream that waits for it AFTER this call has returned sees the bucket complete (the
 * gradient exchange of trainer.py:59-70 DDP then overlaps the rest of the backward pass).  Asynchronous for device
 * buffers like every other entry. */
typedef struct f5b_grad {
  const char* name;
  float* data; /* fp32, numel elements, same memory space as the inputs */
  int64_t numel;
  void* ready_event; /* cudaEvent_t or NULL */
} f5b_grad;
int f5b_cfm_loss_tail_grads(f5b_plan* plan, const float* x1, const float* x0, const int64_t* text, const float* time,
                            const int32_t* lens, const uint8_t* span, int32_t B, int32_t N, int32_t nt,
                            int32_t drop_audio_cond, int32_t drop_text, float* loss, const f5b_grad* grads,
                            int32_t n_grads, int32_t mem, void* stream);

/* Optimiser step of the training loop over FLAT fp32 device buffers (model/trainer.py:118-131 AdamW with
 * torch.optim.AdamW semantics, :387 clip_grad_norm_, :395 EMA): f5b_op_sumsq gives the squared gradient norm (double,
 * fixed reduction tree; scratch >= 1024 doubles), f5b_op_adamw applies one update, reading the gradient as
 * grad * grad_scale (data-parallel average x clip coefficient) and optionally refreshing an EMA copy.  Asynchronous. */
int f5b_op_sumsq(const float* x, int64_t n, double* scratch, double* out, int32_t device, void* stream);
int f5b_op_adamw(float* param, const float* grad, float* exp_avg, float* exp_avg_sq, float* ema, int64_t n, float lr,
                 float beta1, float beta2, float eps, float weight_decay, int32_t step, float grad_scale,
                 float ema_decay, int32_t device, void* stream);

/* Allocation-free variant of f5b_op_sumsq.  NO counterpart in the reference: added
 * for callers that cannot call cudaMalloc (CUDA-graph capture, tight serving loops, harnesses that forbid
 * it).  The 1024 double partials that f5b_op_sumsq takes as `scratch` live in a file-scope `static __device__` array
 * of the library (8 KB on every GPU that loads it, held until exit), so the call allocates nothing on host or device;
 * x and out are device pointers supplied by the caller.  Limit: n <= 2^31 elements (F5B_ERR_INVALID with a message
 * beyond it).  All calls share the one array: calls that overlap on the GPU (different streams or threads) race; calls
 * ordered on one stream are safe.  Nothing needs clearing between calls (every partial is overwritten).  Asynchronous. */
int f5b_op_sumsq_static(const float* x, int64_t n, double* out, int32_t device, void* stream);

/* Replaces MelSpec.forward (model/modules.py:78-104,133-146): wav [B, nw] -> mel [B, 100, nw/256 + 1]. */
int f5b_mel(const float* wav, int32_t B, int32_t nw, float* mel, int32_t device, int32_t mem, void* stream);

/* Replaces the reference-audio preparation of infer_batch_process (infer/utils_infer.py:523-533): channel mix-down,
 * rms = sqrt(mean(mono^2)), gain = target_rms / rms when rms < target_rms (else 1; target_rms <= 0 disables it), and
 * torchaudio.transforms.Resample(sr, 24000) as a strided FIR: out[f * n_new + j] = gain * sum_k table[j][k] *
 * xpad[f * n_orig + k], xpad = mono zero-padded by `width` in front.  wav [channels, nw]; table [n_new, taps] (the
 * caller's windowed-sinc filter bank, taps = 2 * width + n_orig; {1.0} with n_new = n_orig = taps = 1, width = 0 when
 * the rate already matches); out [out_len]; *rms_host (may be NULL) receives the measured rms. */
int f5b_ref_audio(const float* wav, int32_t channels, int64_t nw, const float* table, int32_t n_new, int32_t taps,
                  int32_t n_orig, int32_t width, float target_rms, float* out, int64_t out_len, float* rms_host,
                  int32_t device, int32_t mem, void* stream);

/* Replaces load_vocoder("vocos") + Vocos.decode (infer/utils_infer.py:114-137,703): mel [B,100,T] -> wav [B,(T-1)*256]. */
int f5b_vocos_create(const f5b_tensor* weights, int32_t n_weights, int32_t mode, int32_t device, f5b_vocos** out);
void f5b_vocos_destroy(f5b_vocos* v);
int f5b_vocos_decode(f5b_vocos* v, const float* mel, int32_t B, int32_t T, float* wav, int32_t mem, void* stream);
/* The same over a batch of utterances of different length (the request level decodes a whole micro-batch in one call):
 * lens [B] int32 (same memory space as mel; NULL = all T): row b holds lens[b] frames, the convolutions' zero padding
 * applies at ITS end; wav [B, (T-1)*256], row b valid up to (lens[b]-1)*256 samples, zeros after. */
int f5b_vocos_decode_varlen(f5b_vocos* v, const float* mel, const int32_t* lens, int32_t B, int32_t T, float* wav,
                            int32_t mem, void* stream);
int f5b_vocos_reserve(f5b_vocos* v, int32_t B, int32_t T);

/* Kernel-level entry points used by the parity tests and the roofline benchmark (device pointers only).
 *   gemm:      out[M,N] = act(A[M,K] W[N,K]^T + bias); mode selects fp32 SIMT or bf16 tcgen05 (inputs fp32, converted)
 *   attention: qkv [batches*rows, 3*heads*64] fp32 -> out [batches*rows, heads*64] fp32 */
int f5b_op_gemm(int32_t mode, const float* A, const float* W, const float* bias, int32_t M, int32_t N, int32_t K,
                int32_t act, float* out, int32_t force_bn, int32_t device, void* stream);
/* istft: the Vocos ISTFTHead after its linear layer (exp, clip 1e2, cos / sin, irFFT 1024, Hann, overlap-add, envelope,
 *            trim): z [B*T, ldz] (cols [0,513) log-magnitude, [513,1026) phase), lens [B] or NULL -> wav [B, (T-1)*256].
 *            Witnesses in the reference: runtime/triton_trtllm/scripts/export_vocoder_to_onnx.py:51-59, conv_stft.py:193-234. */
int f5b_op_istft(const float* z, int64_t ldz, const int32_t* lens, int32_t B, int32_t T, float* wav, int32_t device,
                 void* stream);
int f5b_op_attention(int32_t mode, const float* qkv, const int32_t* kv_len, int32_t batches, int32_t rows,
                     int32_t heads, float* out, int32_t device, void* stream);
/* Backward of the same attention (what autograd runs behind modules.py:499-508 under trainer.py:384): given qkv and
 * d_out [batches*rows, heads*64], dqkv [batches*rows, 3*heads*64] receives dq | dk | dv.  mode fp32: CUDA-core kernels;
 * mode bf16: tcgen05 kernels (forward recompute with log-sum-exp, then tc_attn_dq / tc_attn_dkv). */
int f5b_op_attention_bwd(int32_t mode, const float* qkv, const float* d_out, const int32_t* kv_len, int32_t batches,
                         int32_t rows, int32_t heads, float* dqkv, int32_t device, void* stream);
/* Times `iters` back-to-back launches of one kernel with CUDA events on `stream`; returns average ms in *ms.
 *   which: 0 = bf16 tcgen05 GEMM [M,K]x[N,K] (64/128/256: forced tile width; 4: gate*y+residual epilogue), 1 = bf16 tcgen05 attention (M = rows, N = heads, K = batches),
 *          2 = LN+modulate (M rows, N = dim). */
int f5b_op_time(int32_t which, int32_t M, int32_t N, int32_t K, int32_t iters, float* ms, int32_t device, void* stream);

/* Debug timeline of one attention CTA (clock64 stamps, see csrc/runtime.cu); out = int64[3*64*8]. */
int f5b_debug_attn_trace(int32_t rows, int32_t heads, int32_t batches, int64_t* out, int32_t device);

/* Debug in-situ kernel timeline (library built with F5B_KTRACE=1 only; otherwise F5B_ERR_INVALID).
 *   op 0: start -- allocate room for `capacity` records on `device` and arm the per-step kernels;
 *   op 1: stop  -- copy min(*n, capacity) records of 4 x int64 {t_entry, t_after_dependency_wait, t_exit (globaltimer
 *                  ns), kernel_id << 32 | cta} to `out`, set *n to the number recorded, release the buffer. */
int f5b_debug_ktrace(int32_t op, int64_t capacity, int64_t* out, int64_t* n, int32_t device);

/* Number of kernel launches issued by this library since load (for bench.py's gpu_launches). */
int64_t f5b_launch_count(void);
/* Kernel launches inside the last f5b_sample / f5b_vocos_decode / f5b_mel call (graph nodes count individually). */
int64_t f5b_last_call_launches(void);

#ifdef __cplusplus
}
#endif
#endif /* F5B200_H_ */
