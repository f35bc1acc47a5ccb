#!/bin/bash
# Stage 26: per-kernel durations of the decoder step at the reference example's shape (B=2, H=32, 4096 keys).
mkdir -p gpurun_out
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:decode_ -s 40 -c 16 --csv --log-file gpurun_out/launches_step_2_32_4096.csv \
  python tools/decoder_step_time.py 2 32 4096 4 > gpurun_out/ncu_step.log 2>&1; echo "ncu rc=$?"
grep -o '"decode_[a-z_]*kernel.*' gpurun_out/launches_step_2_32_4096.csv | cut -d, -f1,7- | tail -16
