#!/bin/bash
# N-GPU pass after the transposed decode kernel: C4 bench line and the KV-length sharded decode (8-step regions).
N=${1:-8}
mkdir -p gpurun_out
export NCCL_DEBUG=WARN
run() { timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $1 "${@:3}" > gpurun_out/$2.json 2> gpurun_out/$2.err; echo "$2 rc=$?"; cat gpurun_out/$2.json; }
run 29521 bench2_c4_n$N bench.py --gpus $N --workload c4 --steps 30 --warmup 3 --no-secondary
run 29522 shard2_32k_n$N tools/shard_decode_bench.py 32768 30
run 29523 shard2_256k_n$N tools/shard_decode_bench.py 262144 20
run 29524 shard2_1m_n$N tools/shard_decode_bench.py 1048576 20
