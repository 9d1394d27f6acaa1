#!/usr/bin/env python
"""Refresh the GEMV-phases entry of profiles/r02_ncu_summary.json from an `ncu --set full` capture of the persistent decode
kernel, stamping it with the current commit and the hash of the kernel sources (bench.py only trusts `roofline.traffic`
from a capture whose hash equals the build's).

capture (on the B200 box, after the same command has exited 0 without ncu):
    ncu --set full --clock-control none --import-source on -k regex:decode_mega -s 9 -c 1 -o gpurun_out/r02_mega_lin \
        python tools/mega_steps.py 8 lin
here:
    ncu -i gpurun_out/r02_mega_lin.ncu-rep --page raw --csv > profiles/r02_mega_lin_ncu_raw.csv
    python tools/ncu_summary.py profiles/r02_mega_lin_ncu_raw.csv"""
import csv
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    raw = sys.argv[1]
    rows = list(csv.reader(open(raw)))
    hdr, units, val = rows[0], rows[1], rows[-1]
    col = {n: i for i, n in enumerate(hdr)}

    def get(name, scale=None):
        v = float(val[col[name]].replace(",", ""))
        u = units[col[name]]
        if scale == "bytes":
            v *= {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}[u]
        if scale == "us":
            v *= {"ns": 1e-3, "us": 1, "ms": 1e3}.get(u, 1)
        return v

    from bench import kernel_source_sha16
    commit = subprocess.check_output(["git", "rev-parse", "--short", "HEAD"], cwd=ROOT).decode().strip()
    rd, wr = get("dram__bytes_read.sum", "bytes"), get("dram__bytes_write.sum", "bytes")
    stalls = {}
    for n in hdr:
        if n.startswith("smsp__average_warps_issue_stalled_") and n.endswith("_per_issue_active.ratio") and "not_issued" not in n:
            k = n[len("smsp__average_warps_issue_stalled_"):-len("_per_issue_active.ratio")]
            v = float(val[col[n]].replace(",", "") or 0)
            if v >= 0.05 and k != "selected":
                stalls[k] = round(v, 3)
    path = os.path.join(ROOT, "profiles", "r02_ncu_summary.json")
    summ = json.load(open(path))
    e = summ["decode_mega_kernel_mode1_gemv_phases"]
    e.update({
        "captured_at_commit": commit,
        "kernel_source_sha16": kernel_source_sha16(),
        "gpu__time_duration_us": round(get("gpu__time_duration.sum", "us"), 3),
        "dram__bytes_read_MB": rd / 1e6,
        "dram__bytes_write_MB": wr / 1e6,
        "traffic_bytes_per_launch": rd + wr,
        "smsp__issue_active_pct": get("smsp__issue_active.avg.pct_of_peak_sustained_active") if "smsp__issue_active.avg.pct_of_peak_sustained_active" in col else e.get("smsp__issue_active_pct"),
        "warp_instructions": get("smsp__inst_executed.sum"),
        "stall_per_issue": dict(sorted(stalls.items(), key=lambda kv: -kv[1])),
        "registers_per_thread": get("launch__registers_per_thread"),
        "block_size": int(get("launch__block_size")),
    })
    json.dump(summ, open(path, "w"), indent=1)
    print("updated", path, "commit", commit, "hash", e["kernel_source_sha16"], "traffic MB", (rd + wr) / 1e6, "us",
          e["gpu__time_duration_us"])


if __name__ == "__main__":
    main()
