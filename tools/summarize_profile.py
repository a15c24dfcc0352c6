#!/usr/bin/env python
"""Turn the scratch ncu outputs of a gpurun call into the committed summaries under profiles/.
  python tools/summarize_profile.py <launches.csv> <report.ncu-rep> <tag>
Writes profiles/<tag>_launches.csv (kernel, launches, total us, share), profiles/<tag>_<kernel>.txt
(the ncu metrics the roofline uses) and updates profiles/roofline_traffic.json."""
import collections
import csv
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
launches, rep, tag = sys.argv[1:4]
out = os.path.join(ROOT, "profiles")
os.makedirs(out, exist_ok=True)

rows = list(csv.reader(open(launches)))
h = [i for i, r in enumerate(rows) if r and r[0] == "ID"][0]
H = rows[h]
ki, vi = H.index("Kernel Name"), H.index("Metric Value")
agg = collections.OrderedDict()
seq = []
for r in rows[h + 1:]:
    if len(r) > vi:
        name = r[ki].split("(")[0].replace("void ", "").strip()
        t = float(r[vi].replace(",", "")) / 1e3
        a = agg.setdefault(name, [0, 0.0])
        a[0] += 1; a[1] += t
        seq.append((name, t))
tot = sum(v[1] for v in agg.values())
# one timed step = the launches from one k_morton to the next; report one device-resident 1M-query step
starts = [i for i, (nm, _) in enumerate(seq) if nm.endswith("k_morton")]
# (the longest one: bench.py's later end-to-end steps run the same kernels on 131 072-query chunks)
groups = [seq[a:b] for a, b in zip(starts, starts[1:])]
groups = [g for g in groups if not any("k_split" in nm or "k_partition" in nm for nm, _ in g)]  # not a map build
fast = [g for g in groups if any("k_grid_fast" in nm for nm, _ in g)]  # the C2 steps, not the C1 render's gathers
groups = fast or groups
step = max(groups, key=lambda g: sum(t for _, t in g)) if groups else []
with open(os.path.join(out, f"{tag}_step_launches.csv"), "w") as f:
    f.write("# launches of ONE pm_knn step (k_morton .. last query kernel), in launch order; same ncu pass as the full list\n")
    f.write("kernel,us,share_of_step\n")
    st = sum(t for _, t in step) or 1.0
    for nm, t in step:
        f.write(f"\"{nm[:80]}\",{t:.1f},{t / st:.4f}\n")
with open(os.path.join(out, f"{tag}_launches.csv"), "w") as f:
    f.write("# ncu --metrics gpu__time_duration.sum --clock-control none (cold-cache, serialised launches: compare shares)\n")
    f.write("kernel,launches,total_us,share\n")
    for k, (c, t) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        f.write(f"\"{k}\",{c},{t:.1f},{t / tot:.4f}\n")

raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
R = list(csv.reader(raw.splitlines()))
Hh, U = R[0], R[1]
want = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "dram__throughput.avg.pct_of_peak_sustained_elapsed",
        "lts__t_bytes.sum", "lts__t_sector_hit_rate.pct", "l1tex__t_sector_hit_rate.pct", "l1tex__t_bytes.sum",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "launch__registers_per_thread", "launch__grid_size", "launch__block_size",
        "smsp__inst_executed.sum", "sm__inst_executed.avg.per_cycle_active", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
        "smsp__thread_inst_executed_per_inst_executed.ratio", "sm__cycles_elapsed.avg", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
        "lts__throughput.avg.pct_of_peak_sustained_elapsed", "l1tex__throughput.avg.pct_of_peak_sustained_elapsed"]
captured = {}  # short kernel name -> metric values of its first captured launch
for V in R[2:]:
    if len(V) != len(Hh):
        continue
    kname = V[Hh.index("Kernel Name")] if "Kernel Name" in Hh else "kernel"
    short = kname.split("(")[0].split("::")[-1].split("<")[0]
    if short in captured:
        continue
    vals = {}
    with open(os.path.join(out, f"{tag}_{short}_ncu.txt"), "w") as f:
        f.write(f"# ncu --set full --clock-control none, one launch of {kname}\n")
        for w in want:
            if w in Hh:
                i = Hh.index(w)
                vals[w] = (V[i], U[i])
                f.write(f"{w:70s} {V[i]:>18s} {U[i]}\n")
        for i, hname in enumerate(Hh):
            if hname.startswith("smsp__average_warps_issue_stalled") and hname.endswith("per_issue_active.ratio"):
                try:
                    if float(V[i].replace(",", "")) >= 0.15:
                        f.write(f"{hname:70s} {V[i]:>18s} {U[i]}\n")
                except ValueError:
                    pass
    captured[short] = vals


def to_bytes(v, u):
    x = float(v.replace(",", ""))
    return x * {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}[u]


tr = os.path.join(out, "roofline_traffic.json")
j = json.load(open(tr)) if os.path.exists(tr) else {}
for short, vals in captured.items():
    if "dram__bytes_read.sum" in vals:
        j[f"{short}_k64_dram_bytes_per_launch"] = int(to_bytes(*vals["dram__bytes_read.sum"]) + to_bytes(*vals["dram__bytes_write.sum"]))
        j[f"{short}_source"] = f"profiles/{tag}_{short}_ncu.txt"
json.dump(j, open(tr, "w"), indent=1)
print(open(os.path.join(out, f"{tag}_step_launches.csv")).read())
for short in captured:
    print(open(os.path.join(out, f"{tag}_{short}_ncu.txt")).read())
