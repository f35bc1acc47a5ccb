#!/bin/bash
# Round 2: two-level schedule + head-group host pieces: tests, timing, e2e lines, ncu traffic of the prefill kernel.
mkdir -p gpurun_out
( timeout 1500 python -m pytest tests -m gpu -q 2>&1 | tail -12 ) > gpurun_out/r2_fwd12_pytest.log 2>&1
tail -3 gpurun_out/r2_fwd12_pytest.log
{
for cfg in c3 c5 c3r8 c5r8 c2 c3nc; do timeout 300 python tools/fwd_time.py $cfg 10; done
} > gpurun_out/r2_fwd12_time.log 2>&1
grep fwd_time gpurun_out/r2_fwd12_time.log | sed 's/SM clock median(2nd half) //'
for wl in c3 c5 c2; do
  timeout 600 python bench.py --steps 10 --warmup 3 --workload $wl --no-secondary --no-cpu-baseline > gpurun_out/r2_fwd12_bench_$wl.json 2> gpurun_out/r2_fwd12_bench_$wl.err
  python - <<PY
import json
d = json.loads(open("gpurun_out/r2_fwd12_bench_$wl.json").read().strip().splitlines()[-1])
print("$wl: value %.1f ms %.3f | e2e ms %.2f floor %.2f ratio %.3f | sustained %s" % (d["value"], d["ms_per_step"], d["e2e"]["ms_per_step"], d["e2e"]["copy_floor_ms"], d["e2e"]["over_copy_floor"], d.get("sustained", {}).get("value")))
PY
done
python tools/prof_cmd.py fwd > gpurun_out/r2_ncu_plain_fwd.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:fwd_kernel -s 1 -c 1 -f -o gpurun_out/r2_fwd_prof python tools/prof_cmd.py fwd > gpurun_out/r2_ncu_fwd.log 2>&1
echo "fwd capture rc=$?"
