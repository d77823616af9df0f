"""CPU emulation of where bf16 rounding enters the sampling path (tooling, not product): monkeypatches the oracle's
linear / conv / attention contractions to round operands to bf16 for selected name groups and reports the final-mel
max-abs error against the fp32 oracle.  Used to decide which contractions stay fp32 in F5B_MODE_BF16."""
import sys, os, re
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch, torch.nn.functional as F
from oracle import f5_oracle as O

def r(x): return x.to(torch.bfloat16).to(torch.float32)

GROUPS = {
    "in_proj": r"^input_embed\.proj$",
    "proj_out": r"^proj_out$",
    "qkv": r"\.to_[qkv]$",
    "attn_out": r"\.to_out\.0$",
    "ff": r"\.ff\.",
    "text": r"^text_embed",
}
active = set()
orig_lin = O._lin
def lin(sd, name, x):
    for g in active:
        if g in GROUPS and re.search(GROUPS[g], name):
            return F.linear(r(x), r(sd[name + ".weight"]), sd.get(name + ".bias"))
    return orig_lin(sd, name, x)
O._lin = lin
orig_matmul = torch.matmul
def mm(a, b):
    if "attn_core" in active and a.dim() == 4:
        return orig_matmul(r(a), r(b))
    return orig_matmul(a, b)
orig_conv = F.conv1d
def conv(x, w, b=None, **k):
    if "conv_pos" in active and w.shape[-1] == 31:
        return orig_conv(r(x), r(w), b, **k)
    return orig_conv(x, w, b, **k)

def run(cfg, sd, cond, text, N, steps):
    torch.matmul = mm; F.conv1d = conv
    try:
        return O.cfm_sample(sd, cfg, cond, text, N, steps=steps, cfg_strength=2.0, sway_sampling_coef=-1.0, seed=1)
    finally:
        torch.matmul = orig_matmul; F.conv1d = orig_conv

if __name__ == "__main__":
    depth = int(sys.argv[1]) if len(sys.argv) > 1 else 3
    N, steps = 300, 6
    cfg = O.ArchCfg.named("F5TTS_v1_Base", depth=depth)
    sd = O.synth_backbone_weights(cfg, seed=21)
    cond, text = O.synth_inputs(N, N // 3, 40, seed=8)
    torch.set_num_threads(os.cpu_count())
    active.clear()
    ref = run(cfg, sd, cond, text, N, steps)
    print("ref |out| max", float(ref.out.abs().max()), "std", float(ref.out.std()))
    allg = list(GROUPS) + ["attn_core", "conv_pos"]
    for g in [[x] for x in allg] + [allg, [x for x in allg if x not in ("in_proj", "proj_out")]]:
        active.clear(); active.update(g)
        o = run(cfg, sd, cond, text, N, steps)
        e = (o.out - ref.out).abs()
        print(f"{'+'.join(g):60s} max {float(e.max()):.4e} mean {float(e.mean()):.4e}")
