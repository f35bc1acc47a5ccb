"""Turns the three `ncu --set full` captures of tools/gpu_r2_evidence.sh (gpurun_out/r2_{fwd,fwd64,decode}_prof.ncu-rep)
into the committed summaries: profiles/r2_<name>_details.txt (ncu --page details), profiles/r2_<name>_raw_metrics.csv
(the pipe / stall / DRAM rows of --page raw) and profiles/r2_traffic.json (DRAM bytes per launch with the kernel name
and a hash of the kernel's sources; bench.py prints roofline.traffic only while the hash matches).  Runs here (no GPU).

    python tools/ncu_to_profiles.py
"""
import csv
import io
import json
import os
import re
import subprocess
import sys

ROOT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..")
sys.path.insert(0, ROOT)
import bench  # noqa: E402

CAPTURES = {  # workload -> (capture stem, sources the traffic figure depends on, prof_cmd.py argument)
    "c3": ("fwd", ["fwd_sm100.cu", "sm100_ptx.cuh"], "fwd"),
    "c2": ("fwd64", ["fwd_sm100.cu", "sm100_ptx.cuh"], "fwd64"),
    "c4": ("decode", ["decode_sm100.cu", "merge.cuh", "rope.cuh"], "decode"),
}
KEEP = re.compile(r"sm__pipe_tensor|sm__inst_executed_pipe|smsp__pcsamp|dram__bytes|gpu__time_duration|lts__t_sector_hit_rate|"
                  r"sm__cycles_elapsed|smsp__inst_executed.sum|sm__throughput|dram__throughput|launch__")


def ncu(*args):
    return subprocess.run(["ncu", *args], capture_output=True, text=True, check=True).stdout


def to_bytes(value, unit):
    scale = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "Tbyte": 1e12}[unit]
    return float(value.replace(",", "")) * scale


def main():
    traffic = {}
    for wl, (stem, sources, arg) in CAPTURES.items():
        rep = os.path.join(ROOT, "gpurun_out", f"r2_{stem}_prof.ncu-rep")
        if not os.path.exists(rep):
            print("missing", rep)
            continue
        with open(os.path.join(ROOT, "profiles", f"r2_{stem}_details.txt"), "w") as f:
            f.write(ncu("-i", rep, "--page", "details"))
        raw = ncu("-i", rep, "--page", "raw", "--csv")
        rows = list(csv.reader(io.StringIO(raw)))
        names, units, vals = rows[0], rows[1], rows[2]
        col = {n: i for i, n in enumerate(names)}
        with open(os.path.join(ROOT, "profiles", f"r2_{stem}_raw_metrics.csv"), "w") as f:
            f.write(f"# kernel: {vals[col['Kernel Name']]}\n# metric,unit,value  (ncu --set full --clock-control none, one launch; --page raw)\n")
            for n in sorted(names):
                if KEEP.search(n):
                    f.write(f"{n},{units[col[n]]},{vals[col[n]]}\n")
        rd = to_bytes(vals[col["dram__bytes_read.sum"]], units[col["dram__bytes_read.sum"]])
        wr = to_bytes(vals[col["dram__bytes_write.sum"]], units[col["dram__bytes_write.sum"]])
        dur = vals[col["gpu__time_duration.sum"]].replace(",", "")
        dur_ms = float(dur) * {"ns": 1e-6, "us": 1e-3, "ms": 1.0, "s": 1e3}[units[col["gpu__time_duration.sum"]]]
        traffic[wl] = {"bytes_per_launch": rd + wr, "read_bytes": rd, "write_bytes": wr, "kernel": vals[col["Kernel Name"]],
                       "duration_ms_under_ncu": dur_ms,
                       "capture": f"profiles/r2_{stem}_details.txt (ncu --set full --clock-control none, one launch of tools/prof_cmd.py {arg})",
                       "sources": sources, "source_hash": bench.source_hash(*sources)}
        print(wl, f"{(rd + wr) / 1e9:.3f} GB per launch, {dur_ms:.4f} ms under ncu")
    if traffic:
        path = os.path.join(ROOT, "profiles", "r2_traffic.json")
        old = json.load(open(path)) if os.path.exists(path) else {}
        old.update(traffic)
        json.dump(old, open(path, "w"), indent=1)


if __name__ == "__main__":
    main()
