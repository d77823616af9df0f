"""Print the in-kernel timeline of one tcgen05 attention CTA (run under gpurun): python tools/attn_trace.py [rows] [variant]
softmax tile events: 0 wait S, 1 S ready, 2 S in registers, 3 max (+rescale) done, 4 turn acquired, 5 exp + pack + P store
issued, 6 P store complete, 7 P handed over.  MMA warp: 0/3 wait P0/P1, 1/4 seen, 2/5 P*V (+S) issued, 6 kv slot committed."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "korean-f5-tts_b200"))
import torch
from f5b200 import _lib as L
rows = int(sys.argv[1]) if len(sys.argv) > 1 else 940
if len(sys.argv) > 2:
    os.environ["F5B_ATTN_VARIANT"] = sys.argv[2]
buf = torch.zeros(3 * 64 * 8, dtype=torch.int64)
L.check(L.lib().f5b_debug_attn_trace(rows, 16, 2, buf.data_ptr(), 0))
t = buf.view(3, 64, 8)
nkv = (rows + 127) // 128
t0 = int(t[0, 0, 0])
print(f"rows {rows} variant {os.environ.get('F5B_ATTN_VARIANT', 'default')}  (cycles since tile 0's first wait)")
for who, name in ((0, "tile0"), (1, "tile1"), (2, "mma  ")):
    for j in range(nkv):
        print(name, j, " ".join(f"{int(t[who, j, e]) - t0:7d}" if int(t[who, j, e]) else "      -" for e in range(8)))
