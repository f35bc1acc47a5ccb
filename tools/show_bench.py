"""One-screen view of a bench.py JSON line (headline, e2e, roofline, sustained, secondaries).

    python tools/show_bench.py gpurun_out/r2_bench_c3.json [...]
"""
import json
import sys


def show(path):
    with open(path) as f:
        d = json.loads(f.read().strip().splitlines()[-1])
    if "unavailable" in d:
        print(path, d)
        return
    r, e = d.get("roofline") or {}, d.get("e2e") or {}
    print(f"{path}: {d.get('config', {}).get('workload')} n_gpus={d.get('n_gpus')} {d.get('metric')} = {d.get('value'):.1f} {d.get('unit')} "
          f"ms/step {d.get('ms_per_step'):.3f}  frac {r.get('frac')}  traffic {'yes' if r.get('traffic') else 'null'}  clocks {d.get('clocks')}")
    if e:
        print(f"   e2e {e.get('value'):.1f} {e.get('unit')}  ms {e.get('ms_per_step')}  copy floor {e.get('copy_floor_ms')}  over floor {e.get('over_copy_floor')}")
    if d.get("sustained"):
        print("   sustained", {k: d["sustained"].get(k) for k in ("value", "frac", "seconds", "sm_mhz")})
    if d.get("cpu_baseline"):
        print("   cpu_baseline", {k: d["cpu_baseline"].get(k) for k in ("value", "unit", "cores", "kind")})
    for k, v in (d.get("secondary") or {}).items():
        if "error" in v:
            print(f"   {k}: ERROR {v['error']}")
            continue
        clk = v.get("clocks") or {}
        print(f"   {k}: {v.get('value'):.1f} {v.get('unit')}  ms {v.get('ms'):.4f}  frac {v.get('roofline', {}).get('frac'):.3f}  sm_mhz {clk.get('sm_mhz')} {clk.get('reasons')}")


if __name__ == "__main__":
    for p in sys.argv[1:]:
        show(p)
