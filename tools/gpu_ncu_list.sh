#!/bin/bash
# ncu launch list (gpu__time_duration per kernel launch) of one bench step; bench must have exited 0 without ncu first.
mkdir -p gpurun_out
timeout 600 python bench.py --steps 2 --warmup 3 --no-cpu > gpurun_out/bench_pre_ncu.log 2>&1 || { echo "bench failed"; tail -5 gpurun_out/bench_pre_ncu.log; exit 1; }
tail -c 600 gpurun_out/bench_pre_ncu.log
timeout 1500 ncu --metrics gpu__time_duration.sum --clock-control none ${NCU_EXTRA} --launch-skip ${SKIP:-2800} -c ${COUNT:-2900} --csv \
  --log-file gpurun_out/launches.csv python bench.py --steps 1 --warmup 3 --no-cpu > gpurun_out/ncu_list.log 2>&1
echo "ncu rc=$?"; wc -l gpurun_out/launches.csv
