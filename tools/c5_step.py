"""One C5-sized training step on one GPU (BASELINE.json configs[4]: F5TTS_v1_Base, 38 400 frames per GPU), timed.
    python tools/c5_step.py [batch] [frames] [steps]"""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "korean-f5-tts_b200")]
import torch
import bench, f5b200

B = int(sys.argv[1]) if len(sys.argv) > 1 else 40
N = int(sys.argv[2]) if len(sys.argv) > 2 else 960
steps = int(sys.argv[3]) if len(sys.argv) > 3 else 2
dev = torch.device("cuda", 0)
dit, _ = bench.synth_state(0)
cfm = f5b200.CFM(transformer=dit.to(dev), mel_spec_kwargs=dict(mel_spec_type="vocos")).eval()
ts = f5b200.TrainStep(cfm, lr=7.5e-5)
g = torch.Generator().manual_seed(0)
x1 = (torch.randn(B, N, 100, generator=g) * 2 - 3).clamp(min=-11.5).to(dev)
text = torch.randint(0, 58, (B, 170), generator=g).to(dev)
lens = torch.full((B,), N, device=dev)
for i in range(steps + 1):
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    cfm.transformer.plan()  # (re)build the plan from the updated weights: timed separately
    torch.cuda.synchronize()
    t_plan = time.perf_counter() - t0
    loss = ts.step(x1, text, lens=lens)
    torch.cuda.synchronize()
    dt = time.perf_counter() - t0
    print(f"step {i}: plan rebuild {t_plan * 1e3:.0f} ms; loss {float(loss):.4f} {dt:.3f} s  {B * N / dt:.0f} frames/s  grad_norm {ts.last['grad_norm']:.3f} "
          f"{ts.timings_ms()} mem {torch.cuda.max_memory_allocated() / 2**30:.1f} GiB torch", flush=True)
