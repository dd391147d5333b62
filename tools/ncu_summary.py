#!/usr/bin/env python
"""Summarise ncu output brought back in gpurun_out/ into small text files under profiles/ (development tool).
usage: python tools/ncu_summary.py <tag> [launches.csv] [report.ncu-rep]"""
import collections
import csv
import io
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
tag = sys.argv[1]
launches = sys.argv[2] if len(sys.argv) > 2 else os.path.join(ROOT, "gpurun_out", "launches.csv")
rep = sys.argv[3] if len(sys.argv) > 3 else os.path.join(ROOT, "gpurun_out", "prof_umma.ncu-rep")
out_dir = os.path.join(ROOT, "profiles")

if os.path.exists(launches):
    lines = [l for l in open(launches) if not l.startswith("==")]
    r = csv.reader(io.StringIO("".join(lines)))
    hdr = next(r)
    ki, vi = hdr.index("Kernel Name"), hdr.index("Metric Value")
    agg = collections.OrderedDict()
    for row in r:
        if len(row) <= vi:
            continue
        a = agg.setdefault(row[ki], [0, 0.0])
        a[0] += 1
        a[1] += float(row[vi].replace(",", ""))
    tot = sum(a[1] for a in agg.values())
    with open(os.path.join(out_dir, f"{tag}_launch_summary.txt"), "w") as f:
        f.write(f"# ncu --metrics gpu__time_duration.sum --clock-control none  (cold-cache, serialised: compare SHARES)\n")
        f.write(f"# total {tot / 1e6:.3f} ms over {sum(a[0] for a in agg.values())} launches\n")
        for k, a in sorted(agg.items(), key=lambda x: -x[1][1]):
            f.write(f"{a[0]:5d} launches  {a[1] / tot * 100:6.2f}%  avg {a[1] / a[0] / 1e3:10.1f} us  {k[:160]}\n")
    print(open(os.path.join(out_dir, f"{tag}_launch_summary.txt")).read())

if os.path.exists(rep):
    raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(raw)))
    hdr, units = rows[0], rows[1]
    keys = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
            "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
            "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active",
            "sm__throughput.avg.pct_of_peak_sustained_elapsed", "lts__t_sector_hit_rate.pct",
            "lts__throughput.avg.pct_of_peak_sustained_elapsed", "l1tex__m_xbar2l1tex_read_bytes.sum",
            "sm__warps_active.avg.pct_of_peak_sustained_active", "launch__registers_per_thread",
            "sm__cycles_elapsed.avg.per_second", "launch__grid_size", "launch__block_size",
            "launch__shared_mem_per_block_dynamic", "launch__cluster_dim_x"]
    ix = {k: hdr.index(k) for k in keys if k in hdr}
    ni = hdr.index("Kernel Name")
    scale = {"Gbyte": 1e9, "Mbyte": 1e6, "Kbyte": 1e3, "byte": 1}

    def kind(name: str) -> str:
        for pat, k in (("FwdKeepEpi", "fwd_keep"), ("FwdEpi", "fwd"), ("DzEpi", "dz"), (", 0, 1,", "dH"), (", 1, 1,", "dW"),
                       ("probs_to_dz", "convert"), ("gather_windows", "gather")):
            if pat in name:
                return k
        return name[:30]

    traffic = collections.OrderedDict()
    with open(os.path.join(out_dir, f"{tag}_ncu_full_summary.txt"), "w") as f:
        f.write("# ncu --set full --clock-control none --import-source on (consecutive launches of one step, under the profiler)\n")
        for row in rows[2:]:
            f.write(f"\n[{kind(row[ni])}] {row[ni][:150]}\n")
            for k, i in ix.items():
                f.write(f"    {k} [{units[i]}] = {row[i]}\n")
            rd = float(row[ix["dram__bytes_read.sum"]]) * scale[units[ix["dram__bytes_read.sum"]]]
            wr = float(row[ix["dram__bytes_write.sum"]]) * scale[units[ix["dram__bytes_write.sum"]]]
            key = f"{kind(row[ni])}_dram_bytes_per_step"  # summed over that kernel's launches in the captured step
            traffic[key] = traffic.get(key, 0.0) + rd + wr
    json.dump(traffic, open(os.path.join(out_dir, "traffic.json"), "w"), indent=1)
    print(open(os.path.join(out_dir, f"{tag}_ncu_full_summary.txt")).read())
    print(traffic)
