"""Print the in-kernel timeline of one tcgen05 attention CTA (run under gpurun)."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "korean-f5-tts_b200"))
import torch
from f5b200 import _lib as L
rows = int(sys.argv[1]) if len(sys.argv) > 1 else 940
buf = torch.zeros(3 * 64 * 4, dtype=torch.int64)
L.check(L.lib().f5b_debug_attn_trace(rows, 16, 2, buf.data_ptr(), 0))
t = buf.view(3, 64, 4)
nkv = (rows + 127) // 128
t0 = int(t[0, 0, 0])
print("block | tile0: waitS Sready maxdone Pdone | tile1: ... | mma: waitP0 seenP0 waitP1 seenP1   (cycles since start)")
for j in range(nkv):
    row = [int(t[w, j, e]) - t0 for w in range(3) for e in range(4)]
    print(j, " ".join(f"{x:6d}" for x in row))
