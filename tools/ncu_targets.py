"""Small program for `ncu --set full` captures (tools/gpu_ncu_r2.sh): a few launches of the attention kernel at the C2
(N = 940) and C4 (N = 2801) shapes and of the GEMM at the four DiT block shapes (M = 1880), random operands."""
import ctypes as C
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "korean-f5-tts_b200"))
import torch  # noqa: E402
from f5b200 import _lib as L  # noqa: E402

ms = C.c_float()
what = sys.argv[1] if len(sys.argv) > 1 else "attn"
if what == "attn":
    for rows in (940, 2801):
        L.check(L.lib().f5b_op_time(1, rows, 16, 2, 2, C.byref(ms), 0, 0))
        print("attention", rows, ms.value * 1e3, "us")
else:
    for which, n, k in ((0, 3072, 1024), (4, 1024, 1024), (0, 2048, 1024), (4, 1024, 2048)):
        L.check(L.lib().f5b_op_time(which, 1880, n, k, 2, C.byref(ms), 0, 0))
        print("gemm", n, k, ms.value * 1e3, "us")
