#!/bin/bash
# compare kernel-geometry variants (fxtract_b200/_build/var/*.so) on a short bench
for v in "" $(ls fxtract_b200/_build/var/*.so); do
  echo "== variant ${v:-default}"
  FXGPU_LIB=$v python bench.py --records 20000000 --steps 10 --warmup 3 --no-e2e --no-cpu-baseline ${BENCH_EXTRA} | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('value %.0f  scan %.0f GB/s (%.3f)  kernel_ms %.3f pipeline_ms %.3f' % (d['value'], d['roofline']['achieved'], d['roofline']['frac'], d['roofline']['kernel_ms'], d['roofline']['pipeline_ms']))"
done
