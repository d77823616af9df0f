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
if what == "attn_bwd":  # the training step's attention backward at the C5 sequence length (4 of the 40 utterances)
    B, N, H = 4, 960, 16
    g = torch.Generator().manual_seed(0)
    qkv = torch.randn(B * N, 3 * H * 64, generator=g).cuda()
    d_o = torch.randn(B * N, H * 64, generator=g).cuda()
    out = torch.empty(B * N, 3 * H * 64, device="cuda")
    for _ in range(2):
        L.check(L.lib().f5b_op_attention_bwd(1, qkv.data_ptr(), d_o.data_ptr(), None, B, N, H, out.data_ptr(), 0, 0))
    print("attention backward", B, N, H, float(out.abs().max()))
elif what == "gemm_big":  # the CTA-pair tiles at the training step's row count (M = 38 400) and at a C3 batch (32 768)
    for m in (38400, 32768):
        for which, n, k in ((0, 3072, 1024), (0, 2048, 1024), (4, 1024, 2048)):
            L.check(L.lib().f5b_op_time(which, m, n, k, 2, C.byref(ms), 0, 0))
            print("gemm", m, n, k, ms.value * 1e3, "us", 2.0 * m * n * k / ms.value / 1e9, "TFLOP/s")
elif what == "attn":
    for rows in (940, 2801):
        L.check(L.lib().f5b_op_time(1, rows, 16, 2, 2, C.byref(ms), 0, 0))
        print("attention", rows, ms.value * 1e3, "us")
else:
    for which, n, k in ((0, 3072, 1024), (4, 1024, 1024), (0, 2048, 1024), (4, 1024, 2048)):
        L.check(L.lib().f5b_op_time(which, 1880, n, k, 2, C.byref(ms), 0, 0))
        print("gemm", n, k, ms.value * 1e3, "us")
