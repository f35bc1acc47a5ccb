#!/bin/bash
# Multi-GPU pass (gpurun --gpus N -- 'bash tools/gpu_r2_multi.sh N'): the 2-process GPU tests, the sharded bench
# lines for C3 / C5 / C4, the copy ceiling of the box, the KV-length-sharded decode.
N=${1:-2}
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29533"
( timeout 900 python -m pytest tests/test_multi_gpu_gpu.py -q 2>&1 | tail -15 ) > gpurun_out/r2_multi_n${N}_pytest.log 2>&1
tail -3 gpurun_out/r2_multi_n${N}_pytest.log
timeout 600 $TR tools/pcie_ceiling.py > gpurun_out/r2_pcie_n${N}.json 2> gpurun_out/r2_pcie_n${N}.err; tail -c 600 gpurun_out/r2_pcie_n${N}.json
for wl in c3 c5 c4; do
  timeout 900 $TR bench.py --gpus $N --steps 10 --warmup 3 --workload $wl > gpurun_out/r2_bench_${wl}_n${N}.json 2> gpurun_out/r2_bench_${wl}_n${N}.err
  python - <<PY
import json
try:
    d = json.loads(open("gpurun_out/r2_bench_${wl}_n${N}.json").read().strip().splitlines()[-1])
    print("${wl} n=${N}: value %.1f %s ms %.3f | e2e %.1f ms %.2f floor %.2f" % (d["value"], d["unit"], d["ms_per_step"], d["e2e"]["value"], d["e2e"]["ms_per_step"], d["e2e"]["copy_floor_ms"]))
except Exception as e:
    print("${wl}: no line", e); print(open("gpurun_out/r2_bench_${wl}_n${N}.err").read()[-1500:])
PY
done
timeout 600 $TR tools/shard_decode_bench.py > gpurun_out/r2_shard_n${N}.log 2>&1; tail -5 gpurun_out/r2_shard_n${N}.log
