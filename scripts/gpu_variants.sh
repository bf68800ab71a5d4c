#!/bin/bash
# compare kernel variants (fxtract_b200/_build/var/*.so) with the default build on a short bench, interleaved twice
# usage: WORKLOADS="c2 c5c" bash scripts/gpu_variants.sh
mkdir -p gpurun_out
: > gpurun_out/variants.txt
for rep in 1 2; do
for w in ${WORKLOADS:-c2}; do
for v in "" $(ls fxtract_b200/_build/var/*.so 2>/dev/null); do
  FXGPU_LIB=$v python bench.py --workload $w --records ${RECORDS:-20000000} --steps ${STEPS:-10} --warmup 3 --also= --no-e2e --no-cpu-baseline --no-cli --no-verify 2> gpurun_out/variants.err | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); r=d['roofline']
print('%-4s %-28s value %.0f  scan %.0f GB/s (%.3f)  kernel_ms %.3f pipeline_ms %.3f' % ('$w', '${v:-default}'.split('/')[-1], d['value'], r['achieved'], r['frac'], r['kernel_ms'], r['pipeline_ms']))" >> gpurun_out/variants.txt
done
done
done
cat gpurun_out/variants.txt
