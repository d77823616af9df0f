#!/bin/bash
# Source-level (SASS) stall sampling of one kernel inside a real bench step, at the finest sampling interval.
# usage: K=tc_gemm_kernel SKIP=700 CNT=4 bash tools/gpu_ncu_src.sh      (bench must have exited 0 without ncu first)
mkdir -p gpurun_out
timeout 600 python bench.py --steps 2 --warmup 3 --no-cpu > gpurun_out/bench_pre_src.log 2>&1 || { echo "bench failed"; tail -5 gpurun_out/bench_pre_src.log; exit 1; }
K=${K:-tc_gemm_kernel}
timeout 900 ncu --section SourceCounters --section WarpStateStats --section SpeedOfLight --warp-sampling-interval 0 --clock-control none \
  --import-source on -k regex:$K --launch-skip ${SKIP:-700} -c ${CNT:-4} -f -o gpurun_out/src_$K \
  python bench.py --steps 1 --warmup 3 --no-cpu > gpurun_out/ncu_src_$K.log 2>&1
echo "ncu rc=$?"; ls -la gpurun_out/src_$K.ncu-rep
