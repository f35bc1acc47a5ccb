#!/bin/bash
# Round-2 evidence on one GPU: bench lines (C3 default, C4, C2, C5, reference arm), then the ncu passes:
# launch list of the bench command and --set full captures of the prefill and decode kernels.
mkdir -p gpurun_out
timeout 900 python bench.py --steps 20 --warmup 5 > gpurun_out/r2_bench_c3.json 2> gpurun_out/r2_bench_c3.err
timeout 600 python bench.py --steps 20 --warmup 5 --workload c4 --no-secondary > gpurun_out/r2_bench_c4.json 2> gpurun_out/r2_bench_c4.err
timeout 600 python bench.py --steps 20 --warmup 5 --workload c2 --no-secondary > gpurun_out/r2_bench_c2.json 2> gpurun_out/r2_bench_c2.err
timeout 600 python bench.py --steps 5 --warmup 3 --workload c5 --no-secondary > gpurun_out/r2_bench_c5.json 2> gpurun_out/r2_bench_c5.err
timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/r2_bench_ref_c3.json 2> gpurun_out/r2_bench_ref_c3.err
for f in c3 c4 c2 c5 ref_c3; do head -c 400 gpurun_out/r2_bench_$f.json; echo; done
CMD="python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-secondary --no-sustained"
$CMD > gpurun_out/r2_launches_plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r2_launches_bench_c3.csv $CMD > gpurun_out/r2_launches_ncu.log 2>&1
echo "launch list rc=$?"
python tools/prof_cmd.py fwd > gpurun_out/r2_ncu_plain_fwd.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:fwd_kernel -s 1 -c 1 -f -o gpurun_out/r2_fwd_prof python tools/prof_cmd.py fwd > gpurun_out/r2_ncu_fwd.log 2>&1
echo "fwd capture rc=$?"
python tools/prof_cmd.py fwd64 > gpurun_out/r2_ncu_plain_fwd64.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:fwd_kernel -s 1 -c 1 -f -o gpurun_out/r2_fwd64_prof python tools/prof_cmd.py fwd64 > gpurun_out/r2_ncu_fwd64.log 2>&1
echo "fwd64 capture rc=$?"
python tools/prof_cmd.py decode > gpurun_out/r2_ncu_plain_decode.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:decode_splitkv -s 1 -c 1 -f -o gpurun_out/r2_decode_prof python tools/prof_cmd.py decode > gpurun_out/r2_ncu_decode.log 2>&1
echo "decode capture rc=$?"
ls -la gpurun_out/*.ncu-rep
