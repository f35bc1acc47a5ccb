#!/bin/bash
# Stage 14: transposed-MMA decode variant - parity under SFA_DECODE_SWAP=1, then C4 timing against the default.
mkdir -p gpurun_out
SFA_DECODE_SWAP=1 timeout 400 python -m pytest tests/test_decode_gpu.py tests/test_shard_gpu.py -m gpu -q -x > gpurun_out/t_decode_swap.log 2>&1; rc=$?; echo "decode+shard tests (swap) rc=$rc"; tail -6 gpurun_out/t_decode_swap.log
B="python bench.py --workload c4 --steps 30 --no-secondary --no-cpu-baseline"
P='import sys,json; d=json.loads(sys.stdin.read()); print(sys.argv[1], round(d["value"],1), "GB/s", round(d["ms_per_step"]*1e3,1), "us")'
SFA_DECODE_SWAP=0 timeout 200 $B 2>/dev/null | python -c "$P" "swap=0 ctas=3 waves=4:"
for c in 3 4; do for w in 2 3 4 6; do
  SFA_DECODE_SWAP=1 SFA_DECODE_CTAS=$c SFA_DECODE_WAVES=$w timeout 200 $B 2>/dev/null | python -c "$P" "swap=1 ctas=$c waves=$w:"
done; done
SFA_DECODE_SWAP=1 SFA_DECODE_CTAS=4 timeout 200 python tools/decoder_step_time.py 2>&1 | head -3
