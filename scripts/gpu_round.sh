#!/bin/bash
# One GPU-box visit: parity tests, bench line, ncu launch list + full capture of the scan kernel.
set -x
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -3 gpurun_out/pytest_gpu.log
python bench.py --steps 10 --warmup 3 > gpurun_out/bench_c2.json 2> gpurun_out/bench_c2.err; echo "bench rc=$?"
tail -c 3000 gpurun_out/bench_c2.json
SMALL="--records 8000000 --steps 2 --warmup 1 --no-e2e --no-cpu-baseline"
python bench.py $SMALL > gpurun_out/plain_small.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file gpurun_out/launches.csv python bench.py $SMALL > gpurun_out/ncu_launches.log 2>&1
python bench.py $SMALL > gpurun_out/plain_small2.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:scan_tiles -s 2 -c 2 -f -o gpurun_out/prof_scan python bench.py $SMALL > gpurun_out/ncu_full.log 2>&1
ls -la gpurun_out
