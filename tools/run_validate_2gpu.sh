#!/bin/bash
# One 2-GPU call (not product): the sharded GPU tests, then the 2-GPU bench line with BASELINE configs 3-5 appended.
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_sharded.py -m gpu -q -x > gpurun_out/pytest_gpu_w2.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_gpu_w2.log
tail -6 gpurun_out/pytest_gpu_w2.log
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29517 \
  bench.py --gpus 2 --configs 3,4,5 > gpurun_out/bench_n2.json 2> gpurun_out/bench_n2.err; echo "bench rc=$?"
tail -c 600 gpurun_out/bench_n2.err; cut -c1-700 gpurun_out/bench_n2.json
