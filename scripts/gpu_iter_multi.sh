#!/bin/bash
# iteration on the second-stage matchers: field + raw parity tests, then c3 / c5 / c4 bench lines per variant
python -m pytest tests/test_gpu_fields.py tests/test_gpu_raw.py -m gpu -x -q 2>&1 | tail -3
for w in ${WORKLOADS:-c3 c5c c5v c4}; do echo "#### $w"; BENCH_EXTRA="--workload $w" bash scripts/gpu_variants.sh 2>&1; done
