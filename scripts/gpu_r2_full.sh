#!/bin/bash
# Round-2 full visit: what the driver runs at round end (reference arm, default bench line) + the scale tests.
mkdir -p gpurun_out
nproc > gpurun_out/host.txt; free -g >> gpurun_out/host.txt; df -h /dev/shm >> gpurun_out/host.txt
( time python bench.py --impl reference > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err ) 2> gpurun_out/bench_ref.time
( time python bench.py ${BENCH_ARGS} > gpurun_out/bench_full.json 2> gpurun_out/bench_full.err ) 2> gpurun_out/bench_full.time
tail -3 gpurun_out/bench_full.err | cut -c1-400
cat gpurun_out/bench_ref.time gpurun_out/bench_full.time
if [ "${SCALE-1}" != "" ]; then
  ( time timeout 1500 python -m pytest tests/test_gpu_scale.py -m gpu -q --timeout 1200 > gpurun_out/pytest_scale.log 2>&1 ) 2> gpurun_out/pytest_scale.time
  tail -5 gpurun_out/pytest_scale.log | cut -c1-300; cat gpurun_out/pytest_scale.time
fi
python - <<'PY'
import json
try:
    j = json.loads(open("gpurun_out/bench_full.json").read().strip().splitlines()[-1])
    print("c2 value %.0f frac %.3f e2e %s cli %s" % (j["value"], j["roofline"]["frac"], json.dumps(j.get("e2e"))[:300], json.dumps(j.get("e2e_cli"))[:600]))
    for k, v in j.get("workloads", {}).items():
        print(k, "value %.0f frac %.3f kernel_ms %.3f pipeline_ms %.3f selected %d verified %s" % (v["value"], v["roofline"]["frac"], v["roofline"]["kernel_ms"], v["roofline"]["pipeline_ms"], v["selected_records"], v.get("verified")))
    print("cpu", json.dumps(j.get("cpu_baseline"))[:400])
except Exception as e:
    print("no line", e)
PY
