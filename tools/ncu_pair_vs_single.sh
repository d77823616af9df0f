#!/bin/bash
# ncu --set full of the deep-K GEMM (M=1880, N=1024, K=8192: a long steady-state main loop) on single CTAs and CTA pairs.
mkdir -p gpurun_out
cat > /tmp/_pv.py <<'PY'
import ctypes as C, sys
sys.path.insert(0, "korean-f5-tts_b200")
import torch
from f5b200 import _lib as L
ms = C.c_float()
for which in (128, 1128, 256, 1256):
    L.check(L.lib().f5b_op_time(which, 1880, 1024, 8192, 2, C.byref(ms), 0, torch.cuda.current_stream().cuda_stream))
    print(which, ms.value * 1e3, "us", flush=True)
PY
timeout 300 python /tmp/_pv.py || exit 1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:tc_gemm_kernel -f -o gpurun_out/pair_vs_single python /tmp/_pv.py > gpurun_out/ncu_pair.log 2>&1
echo "ncu rc=$?"; ls -la gpurun_out/pair_vs_single.ncu-rep
