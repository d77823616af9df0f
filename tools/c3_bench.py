"""C3 (BASELINE.json configs[2]): F5TTS_v1_Base, 256 variable-length utterances (3-20 s, padded masks), NFE 32, one
GPU's worth of the request-sharded job.  Not the bench.py headline (that is C2); this measures the same path at batch
sizes where the memory-bound kernels and the GEMMs see large M.  Usage (GPU box):
    python tools/c3_bench.py [--utterances 256] [--nfe 32] [--max-rows 16384] [--world 1 --rank 0]
Prints one JSON line.  Lengths follow SURVEY.md 8d: total seconds ~ U(3, 20) with seed 0, prompt = first 3 s."""
import argparse
import json
import os
import sys
import time

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "korean-f5-tts_b200")]
import bench  # noqa: E402  (synthetic weights of the named config)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--utterances", type=int, default=256)
    ap.add_argument("--nfe", type=int, default=32)
    ap.add_argument("--max-rows", type=int, default=16384)
    ap.add_argument("--max-batch", type=int, default=64)
    ap.add_argument("--world", type=int, default=1)
    ap.add_argument("--rank", type=int, default=0)
    ap.add_argument("--batched-vocos", action="store_true", help="one variable-length Vocos call per micro-batch")
    ap.add_argument("--side-stream", action="store_true", help="run on a torch side stream (as bench.py does)")
    ap.add_argument("--no-vocos", action="store_true")
    a = ap.parse_args()
    import f5b200
    from f5b200 import _lib as L
    from f5b200.sharding import micro_batches, my_shard

    dev = torch.device("cuda", 0)
    g = torch.Generator().manual_seed(0)
    secs = torch.rand(a.utterances, generator=g) * 17.0 + 3.0
    frames = [int(s * 24000) // 256 + 1 for s in secs.tolist()]
    n_ref = 3 * 24000 // 256 + 1
    mine = my_shard(frames, a.rank, a.world)
    groups = micro_batches(mine, frames, max_rows=a.max_rows, max_batch=a.max_batch)

    bench.V1_BASE["attn_mask_enabled"] = True
    dit, voc = bench.synth_state(0)
    cfm = f5b200.CFM(transformer=dit.to(dev), mel_spec_kwargs=dict(mel_spec_type="vocos")).eval()
    voc = voc.to(dev).eval()

    def inputs(grp):
        b, n = len(grp), max(frames[i] for i in grp)
        gg = torch.Generator().manual_seed(17 + grp[0])
        cond = (torch.randn(b, n_ref, 100, generator=gg) * 2.0 - 3.0).clamp(min=-11.5)
        nt = max(int(0.18 * n), 8)
        text = torch.randint(0, 58, (b, nt), generator=gg)
        for r, i in enumerate(grp):
            text[r, int(0.18 * frames[i]):] = -1
        dur = torch.tensor([frames[i] for i in grp])
        y0 = torch.randn(b, n, 100, generator=gg)
        return cond.to(dev), text.to(dev), dur.to(dev), y0.to(dev)

    def run(grp, data):
        cond, text, dur, y0 = data
        b = len(grp)
        mel, _ = cfm.sample(cond=cond, text=text, duration=dur if b > 1 else int(dur[0]), steps=a.nfe, cfg_strength=2.0,
                            sway_sampling_coef=-1.0, y0=y0, return_trajectory=False)
        if a.no_vocos:
            return
        if a.batched_vocos and b > 1:
            glen = torch.tensor([frames[i] - n_ref for i in grp], dtype=torch.int32, device=dev)
            voc.decode(mel[:, n_ref:, :].permute(0, 2, 1).contiguous(), glen)
            return
        for r, i in enumerate(grp):
            voc.decode(mel[r:r + 1, n_ref:frames[i], :].permute(0, 2, 1).contiguous())

    data = [inputs(grp) for grp in groups]
    if a.side_stream:
        torch.cuda.set_stream(torch.cuda.Stream(dev))
    run(groups[-1], data[-1])  # warm-up: library / module load, allocator
    run(groups[0], data[0])
    torch.cuda.synchronize()
    l0 = L.lib().f5b_launch_count()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0 = time.perf_counter()
    e0.record()
    for grp, d in zip(groups, data):
        run(grp, d)
    e1.record()
    torch.cuda.synchronize()
    wall = time.perf_counter() - t0
    ms = e0.elapsed_time(e1)
    gen = sum(frames[i] - n_ref for i in mine)
    flops = sum(bench.flops_per_eval(n=frames[i]) for i in mine) * a.nfe
    padded = sum(len(grp) * max(frames[i] for i in grp) for grp in groups)
    pk = bench.peaks()
    print(json.dumps({
        "workload": f"C3: F5TTS_v1_Base, {len(mine)} of {a.utterances} utterances (rank {a.rank}/{a.world}), "
                    f"{sum(frames[i] for i in mine)} frames, NFE {a.nfe}, CFG 2.0 doubled batch, Vocos, attn_mask_enabled",
        "micro_batches": len(groups), "max_rows": a.max_rows, "padding_overhead": round(padded / sum(frames[i] for i in mine) - 1, 4),
        "graph": os.environ.get("F5B_NO_GRAPH") is None,
        "device_s": round(ms / 1e3, 3), "wall_s": round(wall, 3),
        "generated_frames_per_s": round(gen / (ms / 1e3), 1),
        "rtf": round((ms / 1e3) / (gen * 256 / 24000), 6),
        "tflops_unpadded": round(flops / (ms / 1e3) / 1e12, 1),
        "frac_of_sustained_peak": round(flops / (ms / 1e3) / 1e12 / pk["tf_sus"], 4),
        "gpu_launches": int(L.lib().f5b_launch_count() - l0)}))


if __name__ == "__main__":
    main()
