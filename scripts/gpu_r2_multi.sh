#!/bin/bash
# Multi-GPU visit (gpurun --gpus N): equality of N-GPU output with 1 GPU and the reference, then the N-rank bench line.
# usage: N=2 bash scripts/gpu_r2_multi.sh
N=${N:-2}
mkdir -p gpurun_out
nvidia-smi --query-gpu=index,name --format=csv,noheader > gpurun_out/gpus_n$N.txt
if [ "${TESTS-1}" != "" ]; then
  timeout 1200 python -m pytest tests/test_gpu_scale.py -m gpu -q -k multi_gpu --timeout 1000 > gpurun_out/pytest_multi_n$N.log 2>&1
  tail -3 gpurun_out/pytest_multi_n$N.log
fi
python bench.py --impl reference --gpus $N > gpurun_out/bench_ref_n$N.json 2> gpurun_out/bench_ref_n$N.err
( time python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29533 bench.py --gpus $N ${BENCH_ARGS} > gpurun_out/bench_n$N.json 2> gpurun_out/bench_n$N.err ) 2> gpurun_out/bench_n$N.time
tail -2 gpurun_out/bench_n$N.err | cut -c1-300; cat gpurun_out/bench_n$N.time
python scripts/bench_table.py gpurun_out/bench_n$N.json
