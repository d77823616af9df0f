"""Time the tcgen05 GEMM on the DiT block shapes for every tile width (run under gpurun)."""
import ctypes as C, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "korean-f5-tts_b200"))
import torch
from f5b200 import _lib as L

def t(which, M, N, K, iters=50):
    ms = C.c_float()
    L.check(L.lib().f5b_op_time(which, M, N, K, iters, C.byref(ms), 0, torch.cuda.current_stream().cuda_stream))
    return ms.value

M = int(sys.argv[1]) if len(sys.argv) > 1 else 1880
for name, n, k in [("qkv", 3072, 1024), ("attn_out", 1024, 1024), ("ff1", 2048, 1024), ("ff2", 1024, 2048), ("text_pw1", 1024, 512)]:
    row = []
    for w in (0, 128, 192, 256, 1128, 1192, 1256, 4):
        ms = t(w, M, n, k)
        row.append(f"{w}: {ms*1e3:7.2f}us {2.0*M*n*k/ms/1e9:7.1f}TF")
    print(f"{name:9s} M={M} N={n} K={k} | " + " | ".join(row), flush=True)
print("attention", t(1, M // 2, 16, 2) * 1e3, "us")
print("ln_mod", t(2, M, 1024, 0) * 1e3, "us")
