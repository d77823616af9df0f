#!/bin/bash
# ncu --set full captures of the two tensor-core kernels inside a real bench step (one launch each).
mkdir -p gpurun_out
timeout 600 python bench.py --steps 2 --warmup 3 --no-cpu > gpurun_out/bench_pre_full.log 2>&1 || { echo "bench failed"; tail -5 gpurun_out/bench_pre_full.log; exit 1; }
tail -c 300 gpurun_out/bench_pre_full.log
for k in tc_attn_kernel tc_gemm_kernel; do
timeout 900 ncu --set full --clock-control none --import-source on -k regex:$k --launch-skip ${SKIP:-700} -c ${CNT:-4} -f -o gpurun_out/full_$k \
  python bench.py --steps 1 --warmup 3 --no-cpu > gpurun_out/ncu_full_$k.log 2>&1
echo "ncu $k rc=$?"
done
ls -la gpurun_out
