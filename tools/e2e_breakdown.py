"""Where the end-to-end step (host tensors in, waveform out) spends its time beyond the device-resident step."""
import os, sys, time
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
import torch
import bench
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "korean-f5-tts_b200"))
import f5b200

dev = torch.device("cuda", 0)
dit, voc = bench.synth_state(0)
cfm = f5b200.CFM(transformer=dit.to(dev), mel_spec_kwargs=dict(mel_spec_type="vocos")).eval()
voc = voc.to(dev).eval()
kw = dict(duration=bench.N_TOTAL, steps=bench.NFE, cfg_strength=bench.CFG, sway_sampling_coef=bench.SWAY, return_trajectory=False)
ins = [tuple(t.pin_memory() for t in bench.synth_inputs(i)) for i in range(8)]

def sync():
    torch.cuda.synchronize()
    return time.perf_counter()

for rep in range(2):
    acc = {}
    for i in range(8):
        cond, text, y0 = ins[i]
        t0 = sync()
        c, t, y = cond.to(dev, non_blocking=True), text.to(dev, non_blocking=True), y0.to(dev, non_blocking=True)
        t1 = sync()
        mel, _ = cfm.sample(cond=c, text=t, y0=y, **kw)
        t2 = sync()
        m = mel[:, bench.N_REF:, :].permute(0, 2, 1).contiguous()
        wav = voc.decode(m)
        t3 = sync()
        w = wav.cpu()
        t4 = sync()
        for k, v in (("h2d", t1 - t0), ("sample", t2 - t1), ("vocos", t3 - t2), ("d2h", t4 - t3)):
            acc[k] = acc.get(k, 0.0) + v * 1e3 / 8
    print({k: round(v, 3) for k, v in acc.items()}, "ms per step; total", round(sum(acc.values()), 3))
# inside sample: host preparation vs the C call
import f5b200.cfm as C
orig = C.CFM._ode
tt = []
def timed(self, *a, **k):
    t0 = sync(); r = orig(self, *a, **k); tt.append(sync() - t0); return r
C.CFM._ode = timed
for i in range(8):
    cond, text, y0 = ins[i]
    cfm.sample(cond=cond.to(dev), text=text.to(dev), y0=y0.to(dev), **kw)
print("f5b_sample call (incl. its own copies + sync):", round(sum(tt) / len(tt) * 1e3, 3), "ms")
