#!/bin/bash
for m in 0 1 2 4 8 16 32 63; do
  echo "== mask $m"; F5B_PDL_MASK=$m timeout 300 python -m pytest tests/test_gpu_sample.py -m gpu -q -k "golden_fp32 or bf16_vs" 2>&1 | grep -E "passed|failed|^FAILED" | head -8
done
