#!/bin/bash
# ncu launch lists (durations only) of the multi-pattern / exact workloads
mkdir -p gpurun_out
for w in ${WORKLOADS:-c3 c4 c5c}; do
  A="--workload $w --records 8000000 --steps 2 --warmup 1 --no-e2e --no-cpu-baseline"
  python bench.py $A > gpurun_out/plain_$w.log 2>&1 && \
  ncu --metrics gpu__time_duration.sum --clock-control none -c 300 --csv --log-file gpurun_out/launches_$w.csv python bench.py $A > gpurun_out/ncu_$w.log 2>&1
done
