"""Golden output of the TIMED configuration: the fp32 oracle's full `CFM.sample` of bench.py's C2 workload
(F5TTS_v1_Base, depth 22, N = 940, NFE 16, CFG 2.0, sway -1) on bench.py's own synthetic weights and inputs.

The oracle needs minutes of CPU per utterance, so the result is committed: `bench.py` checks the mel of its last timed
utterance against it (16-bit mode tolerance 2e-2, the north-star bound) and `tests/test_gpu_fullsize.py` checks the full
pipeline (mel and, through the oracle's Vocos, the waveform SNR).  Seeds 0 and 1: bench.py cycles its inputs over them.

Run anywhere (no /root/reference needed -- the oracle itself is pinned by tests/test_oracle_golden.py):
  python oracle/make_golden_c2.py   -> tests/golden/c2_oracle_seed{0,1}.npz   (~2-4 min of CPU each)
"""
import os
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "korean-f5-tts_b200")):
    if p not in sys.path:
        sys.path.insert(0, p)

import bench  # noqa: E402  (workload definition: synth_state / synth_inputs)
from oracle import f5_oracle as O  # noqa: E402


def main():
    torch.set_num_threads(os.cpu_count() or 1)
    dit, _ = bench.synth_state(0)
    state = {k: v.detach().float() for k, v in dit.state_dict().items()}
    cfg = O.ArchCfg.named(bench.MODEL)
    for seed in bench.GOLDEN_SEEDS:
        cond, text, y0 = bench.synth_inputs(seed)
        t0 = time.perf_counter()
        with torch.no_grad():
            ref = O.cfm_sample(state, cfg, cond, text, bench.N_TOTAL, steps=bench.NFE, cfg_strength=bench.CFG,
                               sway_sampling_coef=bench.SWAY, y0=y0)
        dt = time.perf_counter() - t0
        path = os.path.join(ROOT, "tests", "golden", f"c2_oracle_seed{seed}.npz")
        np.savez_compressed(path, mel=ref.out[0].numpy().astype(np.float32))
        print(f"seed {seed}: {dt:.1f} s, |mel| max {float(ref.out.abs().max()):.3f}, {os.path.getsize(path)} bytes", flush=True)


if __name__ == "__main__":
    main()
