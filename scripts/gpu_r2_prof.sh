#!/bin/bash
# Launch lists (durations) and one --set full capture of the scan kernel per workload.
# usage: WORKLOADS="c4 c5c c3" RECORDS=8000000 FULL="c5c c3 c4" bash scripts/gpu_r2_prof.sh
mkdir -p gpurun_out
R=${RECORDS:-8000000}
for w in ${WORKLOADS:-c4 c5c c3}; do
  A="--workload $w --records $R --steps 2 --warmup 1 --also= --no-e2e --no-cpu-baseline --no-cli --no-verify ${BENCH_EXTRA}"
  python bench.py $A > gpurun_out/plain_$w.log 2>&1 && \
  ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_$w.csv python bench.py $A > gpurun_out/ncul_$w.log 2>&1
  tail -1 gpurun_out/plain_$w.log | cut -c1-200
done
for w in ${FULL-c5c c3 c4}; do
  A="--workload $w --records $R --steps 2 --warmup 1 --also= --no-e2e --no-cpu-baseline --no-cli --no-verify ${BENCH_EXTRA}"
  ncu --set full --clock-control none --import-source on -k regex:${KREGEX:-scan_tiles} -s ${SKIP:-2} -c 1 -f -o gpurun_out/prof_$w python bench.py $A > gpurun_out/ncu_$w.log 2>&1
  tail -2 gpurun_out/ncu_$w.log | cut -c1-300
done
ls -la gpurun_out/*.ncu-rep
