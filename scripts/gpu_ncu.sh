#!/bin/bash
# ncu full capture of the scan kernel on a small bench run (after the same command passed without ncu)
SMALL="--records 8000000 --steps 2 --warmup 1 --no-e2e --no-cpu-baseline ${BENCH_EXTRA}"
mkdir -p gpurun_out
python bench.py $SMALL > gpurun_out/plain_small.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:${KERNEL:-scan_tiles} -s ${SKIP:-2} -c ${COUNT:-1} -f -o gpurun_out/prof_scan python bench.py $SMALL > gpurun_out/ncu_full.log 2>&1
tail -3 gpurun_out/plain_small.log; tail -3 gpurun_out/ncu_full.log
