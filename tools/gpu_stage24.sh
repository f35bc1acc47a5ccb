#!/bin/bash
# Stage 24: paged decode with the table entry read one step ahead, paged fused decoder step; ncu of the paged kernel.
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_decode_gpu.py -m gpu -q -x > gpurun_out/t_gpu.log 2>&1; rc=$?
echo "gpu decode tests rc=$rc"; tail -8 gpurun_out/t_gpu.log
[ $rc -eq 124 ] && exit 1
timeout 300 python tools/decode_paged_time.py > gpurun_out/paged_c4.log 2>&1; echo "paged time rc=$?"; cat gpurun_out/paged_c4.log
timeout 300 python tools/decode_paged_time.py 1 262144 > gpurun_out/paged_b1_256k.log 2>&1; echo "paged b1 rc=$?"; cat gpurun_out/paged_b1_256k.log
# dense launches of the tool: 3 + 70 + 1 = 74, then page 16: launch 80 is a paged one
timeout 300 ncu --set full --clock-control none --import-source on -k regex:decode_splitkv -s 80 -c 1 -o gpurun_out/r1g_decode_paged \
  python tools/decode_paged_time.py 8 32768 > gpurun_out/ncu_paged.log 2>&1; echo "ncu rc=$?"; grep -i "decode_splitkv" gpurun_out/ncu_paged.log | tail -2
