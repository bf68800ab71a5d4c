#!/bin/bash
# ncu --set full of the scan kernel for several workloads (one call, after the same command passed without ncu)
# usage: WORKLOADS="c5c c4 c3" RECORDS=8000000 bash scripts/gpu_r2_ncu.sh
mkdir -p gpurun_out
for w in ${WORKLOADS:-c5c c4 c3}; do
  A="--workload $w --records ${RECORDS:-8000000} --steps 2 --warmup 1 --also= --no-e2e --no-cpu-baseline --no-cli --no-verify ${BENCH_EXTRA}"
  python bench.py $A > gpurun_out/plain_$w.log 2>&1 && \
  ncu --set full --clock-control none --import-source on -k regex:scan_tiles -s ${SKIP:-2} -c 1 -f -o gpurun_out/prof_$w python bench.py $A > gpurun_out/ncu_$w.log 2>&1
  tail -2 gpurun_out/plain_$w.log | cut -c1-400; tail -2 gpurun_out/ncu_$w.log | cut -c1-300
done
ls -la gpurun_out/*.ncu-rep
