#!/bin/bash
# One GPU visit: parity tests, smoke, bench (+ optional extra command).  Everything lands in gpurun_out/.
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,memory.total --format=csv > gpurun_out/smi.txt 2>&1
nproc > gpurun_out/nproc.txt
timeout 1500 python -m pytest tests -m gpu -x -q -s > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/pytest_gpu.log
timeout 300 python __graft_entry__.py --smoke > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?" | tee -a gpurun_out/smoke.log
timeout 900 python bench.py --steps 10 --warmup 3 > gpurun_out/bench.log 2>&1; echo "bench rc=$?" | tee -a gpurun_out/bench.log
grep -E "passed|failed|error" gpurun_out/pytest_gpu.log | tail -3; tail -5 gpurun_out/smoke.log; tail -2 gpurun_out/bench.log | cut -c1-1500
if [ -n "${1:-}" ]; then bash -c "$*"; fi
