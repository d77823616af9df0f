"""One C3-sized micro-batch (16 utterances x ~1000 frames, CFG doubled: ~32k rows, key masks on) through `sample()`
with a short grid -- the workload for an ncu launch list at batch sizes where the memory-bound kernels stream from HBM.
    F5B_NO_GRAPH=1 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum ... python tools/c3_one_batch.py"""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "korean-f5-tts_b200")]
import bench  # noqa: E402


def main():
    import f5b200

    steps = int(os.environ.get("STEPS", "2"))
    dev = torch.device("cuda", 0)
    bench.V1_BASE["attn_mask_enabled"] = True
    dit, voc = bench.synth_state(0)
    voc = voc.to(dev).eval()
    cfm = f5b200.CFM(transformer=dit.to(dev), mel_spec_kwargs=dict(mel_spec_type="vocos")).eval()
    g = torch.Generator().manual_seed(0)
    b = 16
    durs = torch.randint(900, 1025, (b,), generator=g)
    durs[0] = 1024
    n = int(durs.max())
    cond = (torch.randn(b, 282, 100, generator=g) * 2 - 3).clamp(min=-11.5).to(dev)
    text = torch.randint(0, 58, (b, 180), generator=g).to(dev)
    y0 = torch.randn(b, n, 100, generator=g).to(dev)
    for _ in range(int(os.environ.get("REPS", "1"))):
        cfm.sample(cond=cond, text=text, duration=durs.to(dev), steps=steps, cfg_strength=2.0, sway_sampling_coef=-1.0,
                   y0=y0, return_trajectory=False)
        # the memory-bound tail of the path at the same batch size: Vocos (dwconv + LN, fused iSTFT / overlap-add) over the
        # generated frames with per-row lengths, and the mel front end over 16 x 20 s of audio
        glen = (durs - 282).to(torch.int32).to(dev)
        mel = (torch.randn(b, 100, n - 282, generator=g) * 2 - 3).clamp(min=-11.5).to(dev)
        voc.decode(mel, glen)
        wav = (torch.randn(b, 20 * 24000, generator=g) * 0.1).to(dev)
        cfm.mel_spec(wav)
    torch.cuda.synchronize()
    print("ok", b, n)


if __name__ == "__main__":
    main()
