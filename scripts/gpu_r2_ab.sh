#!/bin/bash
# A/B of an environment switch on one box: usage VAR=FXG_L2_HINT WORKLOADS="c2 c4" bash scripts/gpu_r2_ab.sh
mkdir -p gpurun_out
: > gpurun_out/ab.txt
for rep in 1 2; do
for w in ${WORKLOADS:-c2 c4}; do
  for v in ${VALUES:-1 0}; do
    env ${VAR:-FXG_L2_HINT}=$v python bench.py --workload $w --records ${RECORDS:-20000000} --steps 5 --warmup 3 --also= --no-e2e --no-cpu-baseline --no-cli --no-verify > gpurun_out/ab_$w_$v.json 2> gpurun_out/ab.err
    python - "$w" "$v" gpurun_out/ab_$w_$v.json <<'PY' >> gpurun_out/ab.txt
import json, sys
j = json.loads(open(sys.argv[3]).read().strip().splitlines()[-1]); r = j["roofline"]
print("%s %s=%s value %.0f kernel_ms %.3f pipeline_ms %.3f frac %.3f" % (sys.argv[1], "var", sys.argv[2], j["value"], r["kernel_ms"], r["pipeline_ms"], r["frac"]))
PY
  done
done
done
cat gpurun_out/ab.txt
