#!/bin/bash
# N-GPU pass (gpurun --gpus N): KV-length sharded decode (peer exchange vs NCCL), then the bench lines at N.
N=${1:-2}
mkdir -p gpurun_out
export NCCL_DEBUG=WARN
timeout 120 python tools/shard_decode_bench.py 262144 20 > gpurun_out/shard_n1.json 2> gpurun_out/shard_n1.err; echo "shard n1 rc=$?"; cat gpurun_out/shard_n1.json; tail -2 gpurun_out/shard_n1.err
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 tools/shard_decode_bench.py 262144 20 > gpurun_out/shard_n$N.json 2> gpurun_out/shard_n$N.err; echo "shard n$N rc=$?"; cat gpurun_out/shard_n$N.json; tail -3 gpurun_out/shard_n$N.err
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29514 tools/shard_decode_bench.py 32768 50 > gpurun_out/shard32k_n$N.json 2> gpurun_out/shard32k_n$N.err; echo "shard 32k n$N rc=$?"; cat gpurun_out/shard32k_n$N.json; tail -3 gpurun_out/shard32k_n$N.err
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29512 tools/shard_decode_bench.py 1048576 20 > gpurun_out/shard1m_n$N.json 2> gpurun_out/shard1m_n$N.err; echo "shard 1M n$N rc=$?"; cat gpurun_out/shard1m_n$N.json; tail -3 gpurun_out/shard1m_n$N.err
timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29513 bench.py --gpus $N --steps 10 --warmup 3 > gpurun_out/bench_c3_n$N.json 2> gpurun_out/bench_c3_n$N.err; echo "bench c3 n$N rc=$?"; cat gpurun_out/bench_c3_n$N.json; tail -2 gpurun_out/bench_c3_n$N.err
timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29515 bench.py --gpus $N --workload c4 --steps 20 --warmup 3 --no-secondary > gpurun_out/bench_c4_n$N.json 2> gpurun_out/bench_c4_n$N.err; echo "bench c4 n$N rc=$?"; cat gpurun_out/bench_c4_n$N.json; tail -2 gpurun_out/bench_c4_n$N.err
timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29516 bench.py --gpus $N --workload c5 --steps 5 --warmup 3 --no-secondary > gpurun_out/bench_c5_n$N.json 2> gpurun_out/bench_c5_n$N.err; echo "bench c5 n$N rc=$?"; cat gpurun_out/bench_c5_n$N.json; tail -2 gpurun_out/bench_c5_n$N.err
