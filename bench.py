"""bench.py -- the hot path's headline benchmark (BASELINE.json: RTF / generated mel-frames per second,
F5TTS_v1_Base, NFE 16, CFG 2.0, sway -1, Vocos decode).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]

A "step" = one utterance of config C2 (batch 1, N = 940 frames of which 282 are the reference prompt, 170 text tokens):
CFM.sample (16 CFG-doubled DiT evaluations + Euler) followed by Vocos decode of the 658 generated frames.
  value : generated mel-frames / s, inputs already resident in HBM, CUDA events on the launching stream
  e2e   : the same through the public API (f5b200.CFM.sample + Vocos.decode) with HOST tensors: pinned H2D of the
          prompt mel / noise / tokens and D2H of the waveform inside the timed region (wall clock, synchronised)
  roofline      : the tcgen05 GEMM kernel (the dominant kernel), each DiT shape timed live with CUDA events
  cpu_baseline  : the CPU oracle port of the reference path on this box's host cores, bounded sample (rank 0, N=1)
Multi-GPU: one process per GPU (torchrun), request-sharded, no data-path collective ("weak" scaling): every rank
runs its own utterance stream; NCCL is used for the barrier and the max-over-ranks reduction of the timing only.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
for _p in (ROOT, os.path.join(ROOT, "korean-f5-tts_b200")):
    if _p not in sys.path:
        sys.path.insert(0, _p)

import torch  # noqa: E402

# ---- workload: BASELINE.json configs[1] ---------------------------------------------------------
MODEL = "F5TTS_v1_Base"
N_TOTAL, N_REF, N_TEXT, NFE, CFG, SWAY = 940, 282, 170, 16, 2.0, -1.0
GEN_FRAMES = N_TOTAL - N_REF  # 658 generated frames = 7.02 s at 24 kHz / hop 256
HOP, SR = 256, 24000
METRIC, UNIT = "generated_mel_frames_per_sec", "frames/s"
DTYPE_DETAIL = ("16-bit tensor-core mode: tcgen05 kind::f16, fp32 accumulate; attention q/k/v/P bf16, linear-layer "
                "operands IEEE fp16 (same rate), residual stream / LN statistics / ODE state fp32")


def flops_per_eval(dim=1024, depth=22, heads=16, ff_mult=2, n=N_TOTAL, text_dim=512, mel=100, branches=2):
    """Algorithmic flops of one CFG-doubled function evaluation (872.1 GF for C2)."""
    m = branches * n
    per_tok_layer = 2 * (4 * dim * dim + 2 * ff_mult * dim * dim) + 4 * n * heads * 64
    conv = 2 * 2 * 31 * (dim // 16) * dim
    io = 2 * (2 * mel + text_dim) * dim + 2 * dim * mel
    return m * (depth * per_tok_layer + conv + io)


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        with open(p) as f:
            d = json.load(f)
        return dict(hbm=d["hbm_gbs"], tf=d["bf16_tflops"], tf_sus=d["bf16_tflops_sustained"], src="measured")
    return dict(hbm=6650.0, tf=1590.0, tf_sus=1400.0, src="fallback")  # B200_PROFILING.md fallback


class ClockSampler:
    """nvidia-smi clocks / throttle reasons DURING the timed region (B200_PROFILING.md recipe)."""
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index: int):
        self.rows, self.proc, self.index = [], None, index

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.Q}",
                                          "--format=csv,noheader,nounits", "-lms", "100"], stdout=subprocess.PIPE,
                                         stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._read, daemon=True).start()
        except OSError:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self) -> dict:
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        sm = sorted(int(r[0]) for r in self.rows if r and r[0].isdigit())
        mx = [int(r[1]) for r in self.rows if len(r) > 1 and r[1].isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = [n for i, n in enumerate(names) if any(len(r) > 2 + i and r[2 + i] == "Active" for r in self.rows)]
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": reasons, "samples": len(sm)}


# ---- synthetic model + inputs (random-init weights of the named config; no checkpoints offline) -----------------
V1_BASE = dict(dim=1024, depth=22, heads=16, ff_mult=2, text_dim=512, text_mask_padding=True, conv_layers=4,
               pe_attn_head=None, text_num_embeds=59)  # configs/F5TTS_v1_Base.yaml:24-37; vocab 58 (+1, utils_infer.py:315)


def synth_state(seed: int = 0):
    """Reference-constructor initialisation, then the zero-initialised parameters (AdaLN linears, proj_out, GRN:
    dit.py:214-224, modules.py:234-235) are re-drawn N(0, 0.02^2) so the velocity field is non-trivial."""
    import f5b200

    torch.manual_seed(seed)
    dit = f5b200.DiT(**V1_BASE, precision="bf16")
    g = torch.Generator().manual_seed(seed + 1)
    with torch.no_grad():
        for name, p in dit.named_parameters():
            if "attn_norm.linear" in name or name.startswith("norm_out.") or name.startswith("proj_out.") or ".grn." in name:
                p.copy_(torch.randn(p.shape, generator=g) * 0.02)
    voc = f5b200.Vocos(precision="bf16")
    with torch.no_grad():
        for name, p in voc.named_parameters():
            if name.endswith("norm.weight"):
                p.fill_(1.0)
            elif name.endswith(".gamma"):
                p.fill_(0.125)
            elif name == "head.out.weight":
                p.copy_(torch.randn(p.shape, generator=g) * 0.02)
    return dit, voc


def synth_inputs(seed: int):
    """Log-mel-like prompt (mean -3, std 2, clamp >= ln 1e-5), uniform token ids, seeded noise (cfm.py:196-201)."""
    import math

    g = torch.Generator().manual_seed(1000 + seed)
    cond = (torch.randn(1, N_REF, 100, generator=g) * 2.0 - 3.0).clamp(min=math.log(1e-5))
    text = torch.randint(0, 58, (1, N_TEXT), generator=g)
    torch.manual_seed(seed)
    y0 = torch.randn(N_TOTAL, 100).unsqueeze(0)
    return cond, text, y0


def time_kernel(which: int, M: int, N: int, K: int, iters: int, stream) -> float:
    import ctypes as C

    from f5b200 import _lib as L

    ms = C.c_float()
    L.check(L.lib().f5b_op_time(which, M, N, K, iters, C.byref(ms), torch.cuda.current_device(), stream.cuda_stream))
    return float(ms.value)


# dram__bytes_read.sum + dram__bytes_write.sum per launch of tc_gemm_kernel from the committed `ncu --set full`
# capture of a C2 step (profiles/r1_ncu_full_tc_gemm_v5.txt, cold caches): ff1, ff2, qkv, attn_out.
NCU_TRAFFIC_MB = {"ff1": 8.08, "ff2": 19.63, "qkv": 10.42, "attn_out": 13.68}


def roofline_gemm(stream, pk) -> dict:
    """Dominant kernel = tc_gemm_kernel (tcgen05 GEMM, ~59 % of the step).  Every DiT block shape of one function
    evaluation is timed live: `f5b_op_time` captures 20 launches into a CUDA graph, rotates the weight matrix over
    > 190 MB of copies so W streams from HBM exactly as in the sampling loop (activations stay L2-resident, as they do
    there), and times the graph with CUDA events on its stream.  achieved = sum(algorithmic flops) / sum(avg time);
    algorithmic flops per launch = 2 M N K.  Peak = measured cuBLAS bf16 burst (kernel timed alone)."""
    M = 2 * N_TOTAL
    shapes = [("qkv", 3072, 1024), ("attn_out", 1024, 1024), ("ff1", 2048, 1024), ("ff2", 1024, 2048)]
    tot_f, tot_ms, per = 0.0, 0.0, {}
    for name, n, k in shapes:
        which = 4 if name in ("attn_out", "ff2") else 0  # 4: gate * y + residual epilogue (TMA reduce-add)
        ms = time_kernel(which, M, n, k, 20, stream)
        fl = 2.0 * M * n * k
        per[name] = {"us": round(ms * 1e3, 2), "tflops": round(fl / ms / 1e9, 1),
                     "alg_MB": round((M * k + n * k) * 2 / 1e6, 2), "ncu_dram_MB": NCU_TRAFFIC_MB[name]}
        tot_f += fl
        tot_ms += ms
    att_ms = time_kernel(1, N_TOTAL, 16, 2, 20, stream)
    att_f = 4.0 * N_TOTAL * N_TOTAL * 64 * 16 * 2
    ln_ms = time_kernel(2, M, 1024, 0, 20, stream)
    ln_bytes = M * 1024 * (4 + 2)
    ach = tot_f / tot_ms / 1e9
    return {"bound": "tensor", "kernel": "tc_gemm_kernel (the 4 DiT block shapes, M=1880, weights streamed from HBM)",
            "achieved": round(ach, 1), "peak": pk["tf"],
            "peak_source": pk["src"] + " MEASURED_PEAKS.json bf16_tflops (burst: kernel timed alone)",
            "unit": "TFLOP/s", "frac": round(ach / pk["tf"], 4),
            "traffic": round(sum(NCU_TRAFFIC_MB.values()) / 4 * 1e6), "traffic_note":
            "mean dram read+write bytes per launch over the 4 shapes (ncu --set full, profiles/r1_ncu_full_tc_gemm_v5.txt)",
            "per_shape": per,
            "other_kernels": {
                "tc_attn_kernel": {"us": round(att_ms * 1e3, 2), "tflops": round(att_f / att_ms / 1e9, 1),
                                   "frac_of_tensor_peak": round(att_f / att_ms / 1e9 / pk["tf"], 4)},
                "ln_mod_reg_kernel": {"us": round(ln_ms * 1e3, 2), "GBps": round(ln_bytes / ln_ms / 1e6, 1),
                                      "frac_of_hbm_peak": round(ln_bytes / ln_ms / 1e6 / pk["hbm"], 4),
                                      "note": "operands are L2-resident at batch 1 (11.5 MB): latency-, not HBM-bound"}}}


# ---- CPU leg: the oracle port of the reference path on the host cores (the only place oracle/ is executed) --------
def cpu_reference_sample(state, vstate, seed: int, threads: int):
    """One bounded sample of the workload on the CPU: ONE of the NFE=16 CFG-doubled DiT evaluations (N=940) plus the
    full Vocos decode of 658 frames; the utterance time is extrapolated as 16 x eval + vocos (stated in `sample`)."""
    from oracle import f5_oracle as O  # checker / CPU baseline only

    torch.set_num_threads(threads)
    cfg = O.ArchCfg.named(MODEL)
    cond, text, y0 = synth_inputs(seed)
    cond_p, text_i, dur, cond_mask, step_cond, mask = O.prepare_sample(cond, text, N_TOTAL, None)
    t = O.time_grid(NFE, SWAY)
    with torch.no_grad():
        t0 = time.perf_counter()
        O.backbone_forward(state, cfg, y0, step_cond, text_i, t[3], mask, cfg_infer=True, cache=O.TextCache())
        t_eval = time.perf_counter() - t0
        mel = (torch.randn(1, 100, GEN_FRAMES) * 2.0 - 3.0).clamp(min=-11.5)
        t0 = time.perf_counter()
        O.vocos_decode(vstate, mel)
        t_voc = time.perf_counter() - t0
    return t_eval, t_voc


def cpu_line(state, vstate, threads: int, repeats: int = 1):
    best = None
    for r in range(repeats):
        te, tv = cpu_reference_sample(state, vstate, r, threads)
        tot = NFE * te + tv
        best = tot if best is None else min(best, tot)
    return {"value": round(GEN_FRAMES / best, 3), "unit": UNIT, "cores": threads, "kind": "port",
            "rtf": round(best / (GEN_FRAMES * HOP / SR), 4),
            "sample": f"1 of {NFE} CFG-doubled DiT evaluations (N={N_TOTAL}, fp32, oracle/f5_oracle.py) + full Vocos "
                      f"decode of {GEN_FRAMES} frames; utterance time extrapolated as {NFE} x eval + vocos"}


def run_reference(args, rank: int, world: int):
    """--impl reference: the reference's CPU implementation of the path (oracle port: /root/reference does not exist on
    the GPU box and its third-party deps are absent) on all host threads; rank 0 only."""
    if rank != 0:
        return
    threads = os.cpu_count() or 1
    dit, voc = synth_state(0)
    state = {k: v.detach().float() for k, v in dit.state_dict().items()}
    vstate = {k: v.detach().float() for k, v in voc.state_dict().items()}
    for _ in range(min(args.warmup, 1)):
        cpu_reference_sample(state, vstate, 99, threads)
    tot = 0.0
    for s in range(args.steps):
        te, tv = cpu_reference_sample(state, vstate, s, threads)
        tot += NFE * te + tv
    per = tot / args.steps
    val = GEN_FRAMES / per
    line = {"metric": METRIC, "value": round(val, 3), "unit": UNIT, "impl": "reference", "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": round(per * 1e3, 2), "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "rtf": round(per / (GEN_FRAMES * HOP / SR), 4),
            "config": workload_config(1),
            "cpu_baseline": {"value": round(val, 3), "unit": UNIT, "cores": threads, "kind": "port",
                             "sample": f"per step: 1 of {NFE} CFG-doubled DiT evaluations + full Vocos decode, "
                                       f"utterance time = {NFE} x eval + vocos"},
            "e2e": {"value": round(val, 3), "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(line), flush=True)


def workload_config(world: int) -> dict:
    return {"workload": f"C2: {MODEL} random-init, batch 1 per GPU, N={N_TOTAL} frames (ref {N_REF} + generated "
                        f"{GEN_FRAMES} = 7.02 s), {N_TEXT} text tokens (Korean allophone vocab 58), NFE={NFE} Euler + "
                        f"EPSS/sway {SWAY}, CFG {CFG} (doubled batch), Vocos decode",
            "utterances_per_step": world, "parallelism": f"request-sharded x{world}, no collective",
            "l2": "per-step working set (0.66 GB bf16 weights streamed every function evaluation) exceeds the 126 MB L2"}


# ---- our arm -------------------------------------------------------------------------------------------------------
def run_ours(args, rank: int, world: int, local_rank: int):
    import torch.distributed as dist

    import f5b200
    from f5b200 import _lib as L

    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device -- libf5b200 has no CPU fallback (use --impl reference for the CPU arm)")
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    pk = peaks()
    dit, voc = synth_state(0)
    cfm = f5b200.CFM(transformer=dit.to(dev), mel_spec_kwargs=dict(mel_spec_type="vocos")).eval()
    voc = voc.to(dev).eval()
    stream = torch.cuda.Stream(dev)
    kw = dict(duration=N_TOTAL, steps=NFE, cfg_strength=CFG, sway_sampling_coef=SWAY, return_trajectory=False)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize(dev)

    def reduce_max(x: float) -> float:
        if world == 1:
            return x
        t = torch.tensor([x], device=dev, dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    n_in = args.warmup + args.steps
    host_in = [synth_inputs(rank * 100003 + i) for i in range(n_in)]
    dev_in = [tuple(t.to(dev) for t in x) for x in host_in]
    pin_in = [tuple(t.pin_memory() for t in x) for x in host_in]

    def step_device(i):
        cond, text, y0 = dev_in[i]
        mel, _ = cfm.sample(cond=cond, text=text, y0=y0, **kw)
        return voc.decode(mel[:, N_REF:, :].permute(0, 2, 1).contiguous())

    def step_e2e(i):
        cond, text, y0 = pin_in[i]
        mel, _ = cfm.sample(cond=cond.to(dev, non_blocking=True), text=text.to(dev, non_blocking=True),
                            y0=y0.to(dev, non_blocking=True), **kw)
        wav = voc.decode(mel[:, N_REF:, :].permute(0, 2, 1).contiguous())
        return wav.cpu()

    with torch.cuda.stream(stream):
        # -------- value: device-resident inputs, CUDA events on the launching stream --------
        for i in range(args.warmup):
            step_device(i)
        barrier()
        clk = ClockSampler(local_rank)
        clk.start()
        l0 = L.lib().f5b_launch_count()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream)
        for i in range(args.steps):
            wav = step_device(args.warmup + i)
        e1.record(stream)
        barrier()
        launches = L.lib().f5b_launch_count() - l0
        clocks = clk.stop()
        ms_total = reduce_max(e0.elapsed_time(e1))
        assert torch.isfinite(wav).all(), "non-finite waveform"
        # -------- e2e: public API, host tensors in, waveform out (wall clock, synchronised both sides) --------
        for i in range(min(args.warmup, 2)):
            step_e2e(i)
        barrier()
        t0 = time.perf_counter()
        for i in range(args.steps):
            w = step_e2e(args.warmup + i)
        torch.cuda.synchronize(dev)
        e2e_s = reduce_max(time.perf_counter() - t0)
        barrier()
        h2d = sum(t.numel() * t.element_size() for t in pin_in[0])
        d2h = w.numel() * w.element_size()
        roof = roofline_gemm(stream, pk) if rank == 0 else None

    if rank != 0:
        return
    ms_step = ms_total / args.steps
    val = world * GEN_FRAMES / (ms_step / 1e3)
    e2e_val = world * GEN_FRAMES * args.steps / e2e_s
    gen_s = GEN_FRAMES * HOP / SR
    step_tf = (NFE * flops_per_eval() + 27.0e6 * GEN_FRAMES) / (ms_step / 1e3) / 1e12
    line = {"metric": METRIC, "value": round(val, 1), "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": round(ms_step, 3), "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "bf16", "dtype_detail": DTYPE_DETAIL, "data": "synthetic",
            "config": workload_config(world),
            "rtf": round(ms_step / 1e3 / gen_s, 5), "rtf_e2e": round(e2e_s / args.steps / gen_s, 5),
            "e2e": {"value": round(e2e_val, 1), "unit": UNIT, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h},
            "gpu_launches": int(launches), "clocks": clocks,
            "step_tflops": {"achieved": round(step_tf, 1), "peak_sustained": pk["tf_sus"],
                            "frac": round(step_tf / pk["tf_sus"], 4),
                            "flops_per_step": NFE * flops_per_eval() + 27.0e6 * GEN_FRAMES},
            "roofline": roof}
    if world == 1 and not args.no_cpu:
        state = {k: v.detach().float().cpu() for k, v in dit.state_dict().items()}
        vstate = {k: v.detach().float().cpu() for k, v in voc.state_dict().items()}
        line["cpu_baseline"] = cpu_line(state, vstate, os.cpu_count() or 1)
    print(json.dumps(line), flush=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg (profiling runs)")
    args = ap.parse_args()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if args.impl == "reference":
        run_reference(args, rank, world)
        return
    args.warmup = max(args.warmup, 3)
    if world > 1:
        import torch.distributed as dist

        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        torch.cuda.set_device(local_rank)
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    try:
        run_ours(args, rank, world, local_rank)
    finally:
        if world > 1:
            import torch.distributed as dist

            dist.destroy_process_group()


if __name__ == "__main__":
    main()
