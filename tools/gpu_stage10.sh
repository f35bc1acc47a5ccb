#!/bin/bash
# Stage 10: parallel split merge - decode / shard parity, B=1 long-sequence decode, C4 bench line.
mkdir -p gpurun_out
timeout 400 python -m pytest tests/test_decode_gpu.py tests/test_shard_gpu.py -m gpu -q -x > gpurun_out/t_decode.log 2>&1; echo "decode+shard tests rc=$?"; tail -5 gpurun_out/t_decode.log
timeout 200 python tools/shard_decode_bench.py 262144 20 2>&1 | tail -1
timeout 200 python tools/shard_decode_bench.py 1048576 20 2>&1 | tail -1
timeout 200 python bench.py --workload c4 --steps 20 --no-secondary --no-cpu-baseline 2>/dev/null | tail -1
timeout 200 python tools/decoder_step_time.py 2>&1 | tail -5
