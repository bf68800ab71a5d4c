#!/bin/bash
# Round-2 GPU visit: all GPU parity tests (no -x: every failure is wanted), then short bench lines of each workload.
# usage: RECORDS=20000000 WORKLOADS="c2 c3 c4 c5v c5c" TESTS="tests" bash scripts/gpu_r2_iter.sh
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv,noheader > gpurun_out/gpu.txt 2>&1
if [ "${TESTS-all}" != "" ]; then
  : > gpurun_out/pytest_gpu.log
  for f in ${TESTS:-tests/test_gpu_fused.py tests/test_gpu_raw.py tests/test_gpu_fields.py tests/test_cli_golden.py}; do
    # one process per file: a faulting kernel poisons only its own file
    timeout 1200 python -m pytest $f -m gpu -q --timeout 600 > gpurun_out/pytest_$(basename $f .py).log 2>&1
    echo "$f rc=$?" >> gpurun_out/pytest_gpu.log
    tail -4 gpurun_out/pytest_$(basename $f .py).log >> gpurun_out/pytest_gpu.log
    grep -E "^(FAILED|ERROR)" gpurun_out/pytest_$(basename $f .py).log | cut -c1-200 >> gpurun_out/pytest_gpu.log
  done
  cat gpurun_out/pytest_gpu.log
fi
: > gpurun_out/workloads.txt
for w in ${WORKLOADS:-c2 c3 c4 c5v c5c}; do
  timeout 600 python bench.py --workload $w --records ${RECORDS:-20000000} --steps ${STEPS:-5} --warmup 3 --also '' --no-e2e --no-cpu-baseline --no-cli ${BENCH_EXTRA} \
     > gpurun_out/wl_$w.json 2> gpurun_out/wl_$w.err; echo "$w rc=$?" >> gpurun_out/workloads.txt
  python - "$w" <<'PY' >> gpurun_out/workloads.txt
import json, sys
w = sys.argv[1]
try:
    j = json.loads(open("gpurun_out/wl_%s.json" % w).read().strip().splitlines()[-1])
    r = j["roofline"]
    print("%s value %.0f GB/s  ms/step %.3f  kernel_ms %.3f pipeline_ms %.3f  roofline %.0f GB/s frac %.3f  selected %s launches %s" %
          (w, j["value"], j["ms_per_step"], r["kernel_ms"], r["pipeline_ms"], r["achieved"], r["frac"],
           j["config"].get("selected_records"), j.get("gpu_launches")))
except Exception as e:
    print(w, "no line:", e)
PY
  tail -2 gpurun_out/wl_$w.err | cut -c1-300 >> gpurun_out/workloads.txt
done
cat gpurun_out/workloads.txt
