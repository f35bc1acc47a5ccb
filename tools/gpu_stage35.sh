#!/bin/bash
# Stage 35: programmatic dependent launch on every decode kernel (level 2) against the combine only (level 1).
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q -x > gpurun_out/t_gpu.log 2>&1; rc=$?
echo "gpu tests rc=$rc"; tail -3 gpurun_out/t_gpu.log
[ $rc -eq 124 ] && exit 1
for pdl in 2 1; do
echo "== SFA_DECODE_PDL=$pdl"
SFA_DECODE_PDL=$pdl timeout 300 python bench.py --workload c4 --steps 30 --no-secondary --no-cpu-baseline 2>/dev/null | cut -c1-150
for cfg in "2 32 4096 4" "1 8 32768 4" "8 8 32768 4" "1 32 512 4"; do
  echo "-- decoder step $cfg"; SFA_DECODE_PDL=$pdl timeout 200 python tools/decoder_step_time.py $cfg 2>&1 | grep "decoder_step" | grep -v "fused=0"
done
done
