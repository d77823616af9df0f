"""A/B of the attention kernel's variants on the GPU (F5B_ATTN_VARIANT = SEQ*100 + EARLY*10 + POLY, csrc/tc_attn.cu):
parity of each variant against an fp64 softmax reference (incl. key masks, N = 2801 and the rescale path), then the
per-launch time at the C2 / C4 / C3-batch shapes through f5b_op_time (20 launches in a CUDA graph, random operands)."""
import ctypes as C
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "korean-f5-tts_b200"), os.path.join(ROOT, "tests")):
    sys.path.insert(0, p)
import torch  # noqa: E402

from f5b200 import _lib as L  # noqa: E402
import test_gpu_ops as T  # noqa: E402

PEAK = 1668.1
variants = [int(v) for v in sys.argv[1:]] or [0, 100, 110, 111, 112]
shapes = [("C2", 940, 16, 2), ("C4", 2801, 16, 2), ("C3b", 1024, 16, 32), ("C3s", 512, 16, 32)]
for v in variants:
    os.environ["F5B_ATTN_VARIANT"] = str(v)
    errs = []
    for batches, rows, heads, lens in [(1, 100, 2, None), (2, 333, 3, [333, 201]), (2, 940, 16, None), (1, 2801, 4, None)]:
        g = torch.Generator().manual_seed(rows + 1)
        qkv = torch.randn(batches * rows, 3 * heads * 64, generator=g)
        kv = torch.tensor(lens) if lens else None
        out = T._attention(1, qkv, kv, batches, rows, heads)
        ref = T._attention_ref(qkv.bfloat16().float(), kv, batches, rows, heads)
        errs.append(float((out - ref).abs().max()))
    try:
        T.test_attention_bf16_rising_scores_rescale_path(700, 2)
        rs = "ok"
    except AssertionError as e:
        rs = f"FAIL {e}"
    line = f"variant {v:3d}: parity max-abs {max(errs):.2e} rescale {rs} |"
    for name, rows, heads, batches in shapes:
        ms = C.c_float()
        L.check(L.lib().f5b_op_time(1, rows, heads, batches, 20, C.byref(ms), 0, 0))
        fl = 4.0 * rows * rows * 64 * heads * batches
        tf = fl / ms.value / 1e9
        line += f" {name} {ms.value * 1e3:7.2f} us {tf:6.0f} TF/s ({tf / PEAK:.3f})"
    print(line, flush=True)
