#!/bin/bash
# Stage 8: fused decoder step + graph replay - parity, then timing; full suite last.
mkdir -p gpurun_out
timeout 400 python -m pytest tests/test_decode_gpu.py -m gpu -q -x > gpurun_out/t_decode.log 2>&1; echo "decode tests rc=$?"; tail -8 gpurun_out/t_decode.log
timeout 200 python tools/decoder_step_time.py 2>&1 | tail -8
timeout 200 python bench.py --workload c4 --steps 20 --no-secondary 2>/dev/null | tail -1
timeout 400 python -m pytest tests -m gpu -q -x > gpurun_out/t_gpu.log 2>&1; echo "all gpu tests rc=$?"; tail -4 gpurun_out/t_gpu.log
