#!/bin/bash
mkdir -p gpurun_out
timeout 400 python -m pytest tests/test_decode_gpu.py tests/test_shard_gpu.py -m gpu -q -x > gpurun_out/t_decode.log 2>&1; echo "decode+shard tests rc=$?"; tail -3 gpurun_out/t_decode.log
timeout 200 python tools/shard_decode_bench.py 262144 10 2>&1 | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('B=1 256K:', d['peer'])"
for w in 2 3 4 5 6 8; do
  SFA_DECODE_WAVES=$w timeout 200 python bench.py --workload c4 --steps 30 --no-secondary --no-cpu-baseline 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('waves $w:', round(d['value'],1), 'GB/s', round(d['ms_per_step']*1e3,1), 'us')"
done
