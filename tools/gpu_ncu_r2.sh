#!/bin/bash
# Round-2 profiler evidence, one gpurun call (all ncu): every command first exits 0 without ncu.
#   1. launch list (+ DRAM bytes) of one C2 bench step           -> gpurun_out/r2_launches_c2.csv
#   2. launch list (+ DRAM bytes) of one C3-sized micro-batch     -> gpurun_out/r2_launches_c3_batch.csv
#   3. ncu --set full of tc_attn_kernel (N = 940, 2801) and of tc_gemm_kernel (4 block shapes, M = 1880)
mkdir -p gpurun_out
M="gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum"
timeout 600 python bench.py --steps 1 --warmup 3 --no-cpu --no-c3 --no-c4 > gpurun_out/r2_pre_c2.log 2>&1 &&
timeout 1500 ncu --metrics $M --clock-control none --launch-skip ${SKIP:-8800} -c 2800 --csv --log-file gpurun_out/r2_launches_c2.csv \
  python bench.py --steps 1 --warmup 3 --no-cpu --no-c3 --no-c4 > gpurun_out/r2_ncu_c2.log 2>&1
echo "c2 list rc=$?"
F5B_NO_GRAPH=1 timeout 300 python tools/c3_one_batch.py > gpurun_out/r2_pre_c3.log 2>&1 &&
F5B_NO_GRAPH=1 timeout 1500 ncu --metrics $M --clock-control none -c 4000 --csv --log-file gpurun_out/r2_launches_c3_batch.csv \
  python tools/c3_one_batch.py > gpurun_out/r2_ncu_c3.log 2>&1
echo "c3 list rc=$?"
for w in attn gemm; do
  k=tc_attn_kernel; [ $w = gemm ] && k=tc_gemm_kernel
  timeout 120 python tools/ncu_targets.py $w > gpurun_out/r2_pre_$w.log 2>&1 &&
  timeout 900 ncu --set full --clock-control none --import-source on -k regex:$k -f -o gpurun_out/r2_full_$w \
    python tools/ncu_targets.py $w > gpurun_out/r2_ncu_full_$w.log 2>&1
  echo "full $w rc=$?"
done
ls -la gpurun_out | grep r2_
