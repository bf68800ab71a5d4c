#!/bin/bash
# Round-end visit on one GPU: every GPU test file, then what the driver runs (reference arm + default bench line).
mkdir -p gpurun_out
: > gpurun_out/pytest_final.log
for f in tests/test_gpu_fused.py tests/test_gpu_raw.py tests/test_gpu_fields.py tests/test_cli_golden.py tests/test_gpu_scale.py; do
  timeout 1500 python -m pytest $f -m gpu -q --timeout 1200 > gpurun_out/pytest_$(basename $f .py).log 2>&1
  echo "$f rc=$?" >> gpurun_out/pytest_final.log
  tail -2 gpurun_out/pytest_$(basename $f .py).log >> gpurun_out/pytest_final.log
done
cat gpurun_out/pytest_final.log
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?"; tail -2 gpurun_out/smoke.log
SCALE="" bash scripts/gpu_r2_full.sh
