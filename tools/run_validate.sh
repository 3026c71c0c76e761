#!/bin/bash
# One GPU call (not product): the GPU suite, the N=1 bench line, then the ncu launch list of the same bench command.
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_gpu.log
tail -3 gpurun_out/pytest_gpu.log
timeout 900 python bench.py > gpurun_out/bench_n1.json 2> gpurun_out/bench_n1.err; echo "bench rc=$?"
cut -c1-1500 gpurun_out/bench_n1.json
if [ -n "$NCU" ]; then
  timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r2_launches.csv \
    python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_bench.log 2>&1; echo "ncu rc=$?"
fi
if [ -n "$NCUFULL" ]; then
  timeout 900 ncu --set full --clock-control none --import-source on -k regex:"sparse_update_kernel|pooled_forward_kernel" \
    --launch-skip 4 -c 2 -f -o gpurun_out/r2_fwd_upd_full python tools/tune_update.py X base > gpurun_out/ncu_full_run.log 2>&1
  echo "ncu full rc=$?"; tail -3 gpurun_out/ncu_full_run.log
  ncu -i gpurun_out/r2_fwd_upd_full.ncu-rep --page raw --csv > gpurun_out/r2_fwd_upd_full_raw.csv 2>/dev/null
fi
