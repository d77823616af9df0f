import os, sys
os.environ["F5B_NO_GRAPH"] = "1"; os.environ["F5B_SYNC_DEBUG"] = "1"
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "korean-f5-tts_b200")); sys.path.insert(0, os.path.join(ROOT, "tests"))
import torch
from oracle import f5_oracle as O
from helpers import build_cfm
cfg = O.ArchCfg.named("E2TTS_Base", depth=4)
sd = O.synth_backbone_weights(cfg, seed=21)
N = 260
cond, text = O.synth_inputs(N, N // 3, 40, seed=8)
cfm = build_cfm(cfg, sd, "bf16")
try:
    out, _ = cfm.sample(cond=cond.cuda(), text=text.cuda(), duration=N, steps=2, cfg_strength=2.0, sway_sampling_coef=-1.0, seed=1)
    print("ok", float(out.abs().max()))
except Exception as e:
    print("FAILED:", e)
