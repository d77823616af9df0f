"""Latency of C2-sized requests whose length changes from call to call (serving traffic): with one cached graph every
change of shape re-captures and re-instantiates the ~11 000-node sampling graph; the plan's LRU of graphs replays.
Prints per-shape first-call (capture) and revisit (replay) wall times.  GPU box: python tools/varlen_latency.py"""
import json
import os
import sys
import time

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "korean-f5-tts_b200")]
import bench  # noqa: E402


def main():
    import f5b200

    dev = torch.device("cuda", 0)
    dit, _ = bench.synth_state(0)
    cfm = f5b200.CFM(transformer=dit.to(dev), mel_spec_kwargs=dict(mel_spec_type="vocos")).eval()
    shapes = [940, 700, 820, 512, 640]
    g = torch.Generator().manual_seed(0)
    data = {n: ((torch.randn(1, 282, 100, generator=g) * 2 - 3).to(dev), torch.randint(0, 58, (1, int(0.18 * n)), generator=g).to(dev),
                torch.randn(1, n, 100, generator=g).to(dev)) for n in shapes}

    def call(n):
        cond, text, y0 = data[n]
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        cfm.sample(cond=cond, text=text, y0=y0, duration=n, steps=16, cfg_strength=2.0, sway_sampling_coef=-1.0,
                   return_trajectory=False)
        torch.cuda.synchronize()
        return (time.perf_counter() - t0) * 1e3

    call(shapes[0]); call(shapes[0])  # library warm-up, workspace at its largest
    first = {n: round(call(n), 2) for n in shapes[1:]}
    again = {n: round(min(call(n) for _ in range(1)), 2) for n in shapes[1:]}  # shape changes on every call
    third = {n: round(call(n), 2) for n in shapes[1:]}
    print(json.dumps({"first_call_ms(capture+instantiate+run)": first, "revisit_ms": again, "revisit2_ms": third}))


if __name__ == "__main__":
    main()
