#!/bin/bash
# quick GPU visit: kernel parity tests + short bench lines of the BASELINE configs
python -m pytest tests/test_gpu_raw.py tests/test_gpu_fields.py -m gpu -x -q 2>&1 | tail -8
for w in ${WORKLOADS:-c2 c2c c3 c4 c5v c5c}; do
  echo "== $w"
  python bench.py --workload $w --records ${RECORDS:-20000000} --steps 5 --warmup 3 --no-e2e --no-cpu-baseline 2>&1 | tail -1 | python -c "
import json,sys
try:
    d=json.loads(sys.stdin.read()); r=d['roofline']
    print('value %.0f GB/s  scan %.0f GB/s (%.3f)  kernel_ms %.3f pipeline_ms %.3f selected %d launches %d' % (d['value'], r['achieved'], r['frac'], r['kernel_ms'], r['pipeline_ms'], d['config']['selected_records'], d['gpu_launches']))
except Exception as e: print('FAILED', e)
"
done
