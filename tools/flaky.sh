#!/bin/bash
for i in 1 2 3 4; do
  F5B_NO_PDL=1 timeout 300 python -m pytest tests/test_gpu_sample.py -m gpu -q -s -k "bf16_vs" 2>&1 | grep -E "max-abs|passed|failed"
done
echo "== with PDL"
for i in 1 2; do
  timeout 300 python -m pytest tests/test_gpu_sample.py -m gpu -q -s -k "bf16_vs or golden" 2>&1 | grep -E "max-abs|passed|failed"
done
