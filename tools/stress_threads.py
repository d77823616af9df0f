"""Two host threads sampling on ONE plan while graphs are being re-captured (18 graph keys against an LRU of 8) and each
thread reads its result back after every call.  With `device-sync` each thread ALSO synchronises the whole device, which
CUDA forbids while another thread records (the synchronising thread gets cudaErrorStreamCaptureUnsupported and the
recording is invalidated; the library then runs that call un-captured).
    python tools/stress_threads.py [iterations] [device-sync]"""
import os, sys, threading, traceback
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "korean-f5-tts_b200"), os.path.join(ROOT, "tests")]
import torch
from helpers import O, build_cfm

iters = int(sys.argv[1]) if len(sys.argv) > 1 else 30
DEVICE_SYNC = len(sys.argv) > 2 and sys.argv[2] == "device-sync"
cfg = O.ArchCfg(dim=128, depth=2, heads=2, ff_mult=2, text_dim=64, conv_layers=2)
cfm = build_cfm(cfg, O.synth_backbone_weights(cfg, seed=2), "fp32")
g = torch.Generator().manual_seed(11)
jobs = {}
for n in (72, 100):
    cond, text = O.synth_inputs(n, 20, 7, seed=n)
    jobs[n] = (cond.cuda(), text.cuda(), torch.randn(1, n, 100, generator=g))


def run(n, steps):
    cond, text, y0 = jobs[n]
    out = cfm.sample(cond=cond, text=text, y0=y0, duration=n, steps=steps, cfg_strength=2.0, sway_sampling_coef=-1.0)[0]
    if DEVICE_SYNC:  # forbidden by CUDA while another thread records a graph: the CALLER gets an error
        torch.cuda.synchronize()
    return out.cpu()


run(100, 4)
alone = {(n, s): run(n, s) for n in jobs for s in range(3, 12)}
bad = []


def worker(n):
    try:
        with torch.cuda.stream(torch.cuda.Stream()):
            for it in range(iters):
                s = 3 + (it * (2 if n == 72 else 5)) % 9
                if not torch.equal(run(n, s), alone[(n, s)]):
                    bad.append(f"n={n} steps={s} it={it}: result differs")
    except Exception:
        bad.append(traceback.format_exc())


ths = [threading.Thread(target=worker, args=(n,)) for n in jobs]
[t.start() for t in ths]
[t.join() for t in ths]
print("\n".join(bad) if bad else f"ok: {2 * iters} threaded calls equal their single-threaded results")
sys.exit(1 if bad else 0)
