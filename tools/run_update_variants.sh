#!/bin/bash
# One GPU call (not product): builds of the update kernel timed in one process (tools/tune_update_multi.py),
# then one `ncu --set full` capture of the lookup and the update kernel of the shipped library.
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
args="base=tdfo_b200/libtdfo_b200.so"
for v in "$@"; do args="$args $v=tools/_variants/$v.so"; done
SHAPES=${SHAPES:-config2,1of8_tw} timeout 600 python tools/tune_update_multi.py $args > gpurun_out/tune_multi.log 2>&1; echo "rc=$?" >> gpurun_out/tune_multi.log
cat gpurun_out/tune_multi.log
if [ -n "$NCU" ]; then
  timeout 900 ncu --set full --clock-control none --import-source on -k regex:"sparse_update_kernel|pooled_forward_kernel" \
    --launch-skip 4 -c 2 -f -o gpurun_out/r2_fwd_upd_full python tools/tune_update.py X base > gpurun_out/ncu_run.log 2>&1
  echo "ncu rc=$?"; tail -5 gpurun_out/ncu_run.log
  ncu -i gpurun_out/r2_fwd_upd_full.ncu-rep --page raw --csv > gpurun_out/r2_fwd_upd_full_raw.csv 2>/dev/null
  ls -la gpurun_out/
fi
