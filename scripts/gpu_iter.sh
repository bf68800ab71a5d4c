#!/bin/bash
# one optimisation iteration on the GPU box: raw-path parity tests, variant bench lines, ncu capture of the scan kernel
mkdir -p gpurun_out
python -m pytest tests/test_gpu_raw.py -m gpu -x -q 2>&1 | tail -3
BENCH_EXTRA="--workload ${W:-c2}" bash scripts/gpu_variants.sh 2>&1 | tee gpurun_out/variants.txt
if [ -n "$NCU" ]; then
SMALL="--records 8000000 --steps 2 --warmup 1 --no-e2e --no-cpu-baseline"
ncu --set full --clock-control none --import-source on -k regex:scan_tiles -s 2 -c 1 -f -o gpurun_out/prof_scan python bench.py $SMALL > gpurun_out/ncu_full.log 2>&1
fi
