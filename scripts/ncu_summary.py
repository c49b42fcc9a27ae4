"""Turns gpurun_out ncu artefacts into the small tracked summaries under profiles/.
Usage: ncu_summary.py <tag> [launches.csv] [report.ncu-rep]"""
import collections
import csv
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
tag = sys.argv[1]
launches = sys.argv[2] if len(sys.argv) > 2 else os.path.join(ROOT, "gpurun_out", "launches.csv")
rep = sys.argv[3] if len(sys.argv) > 3 else None
out = []
if launches and os.path.exists(launches):
    rows = list(csv.reader(open(launches)))
    hi = next(i for i, r in enumerate(rows) if "Kernel Name" in r)
    hdr = rows[hi]; ki = hdr.index("Kernel Name"); vi = hdr.index("Metric Value"); ui = hdr.index("Metric Unit")
    agg = collections.defaultdict(lambda: [0, 0.0])
    for r in rows[hi + 1:]:
        if len(r) <= vi:
            continue
        v = float(r[vi].replace(",", ""))
        v = v / 1e3 if r[ui] == "ns" else (v * 1e3 if r[ui] == "ms" else v)
        name = r[ki].split("(")[0].replace("void ", "")
        agg[name][0] += 1; agg[name][1] += v
    tot = sum(v[1] for v in agg.values())
    out.append(f"## ncu launch list ({tag}): gpu__time_duration.sum per kernel, --clock-control none\n")
    out.append(f"total {tot / 1e3:.3f} ms over {sum(v[0] for v in agg.values())} launches (cold-cache, serialised: compare shares)\n")
    out.append("| kernel | launches | total ms | mean us | share |\n|---|---:|---:|---:|---:|")
    for k, v in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        out.append(f"| `{k}` | {v[0]} | {v[1] / 1e3:.3f} | {v[1] / v[0]:.2f} | {v[1] / tot * 100:.1f}% |")
    out.append("")
traffic = {}
if rep and os.path.exists(rep):
    raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(raw.splitlines()))
    hdr, units = rows[0], rows[1]
    want = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
            "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "sm__warps_active.avg.pct_of_peak_sustained_active",
            "launch__registers_per_thread", "l1tex__t_sector_hit_rate.pct", "lts__t_sector_hit_rate.pct",
            "l1tex__throughput.avg.pct_of_peak_sustained_active", "lts__throughput.avg.pct_of_peak_sustained_elapsed",
            "sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active"]
    idx = [hdr.index(w) for w in want]
    ki = hdr.index("Kernel Name")
    out.append(f"## ncu --set full ({tag}), one row per captured launch\n")
    out.append("| kernel | us | DRAM rd MB | DRAM wr MB | DRAM GB/s | dram % | warps act % | regs | L1 hit % | L2 hit % | L1tex % | LTS % | fp64 pipe % |\n|---|---:|---:|---:|---:|---:|---:|---:|---:|---:|---:|---:|---:|")
    seen = collections.Counter()
    for r in rows[2:]:
        name = r[ki].split("(")[0].replace("void ", "")
        seen[name] += 1
        if seen[name] > 2:
            continue
        v = [float(r[i].replace(",", "")) for i in idx]
        mult = lambda i: {"Gbyte": 1e3, "Mbyte": 1.0, "Kbyte": 1e-3, "byte": 1e-6}[units[i]]
        us = v[0] * {"us": 1.0, "ns": 1e-3, "ms": 1e3}[units[idx[0]]]
        rd, wr = v[1] * mult(idx[1]), v[2] * mult(idx[2])
        traffic.setdefault(name.split("<")[0], int((rd + wr) * 1e6))
        out.append(f"| `{name}` | {us:.1f} | {rd:.1f} | {wr:.1f} | {(rd + wr) / us * 1e3:.0f} | {v[3]:.1f} | {v[4]:.1f} | {int(v[5])} | "
                   f"{v[6]:.1f} | {v[7]:.1f} | {v[8]:.1f} | {v[9]:.1f} | {v[10]:.1f} |")
    out.append("")
open(os.path.join(ROOT, "profiles", f"{tag}_ncu_summary.md"), "w").write("\n".join(out) + "\n")
if traffic:
    json.dump(traffic, open(os.path.join(ROOT, "profiles", f"{tag}_traffic.json"), "w"), indent=1)
print("\n".join(out))
