#!/usr/bin/env python
"""Turn the scratch ncu outputs of a gpurun call into the committed summaries under profiles/.
  python tools/summarize_profile.py <tag> <launches.csv> <report.ncu-rep> [<report2.ncu-rep> ...]
Writes profiles/<tag>_launches.csv (kernel, launches, total us, share), profiles/<tag>_step_launches.csv (the launches of
one timed pm_knn step), one profiles/<tag>_<kernel>_ncu.txt per captured kernel (the ncu metrics the roofline uses,
first captured launch; `launches_captured` and the summed DRAM bytes say how many launches the report holds) and
updates profiles/roofline_traffic.json — the per-launch `dram__bytes_read.sum + dram__bytes_write.sum` figures bench.py
prints as `roofline.traffic`:
  knn_step_k64_dram_bytes_per_launch   sum over the kernels of ONE pm_knn step (k_qbin_*, k_knn_fast, k_grid_query, k_kdw)
  k_walk_dram_bytes_per_launch         one stage-1 walk of the bench (4 M paths on the Cornell box)
  map_build_10M_dram_bytes_per_launch  sum over the kernels of one 10 M-photon pm_map_build, if a report holds them"""
import collections
import csv
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
tag, launches = sys.argv[1:3]
reps = sys.argv[3:]
out = os.path.join(ROOT, "profiles")
os.makedirs(out, exist_ok=True)

STEP_FIRST = "k_qbin_count"
STEP_KERNELS = ("k_qbin_count", "k_qbin_scan", "k_qbin_scatter", "k_knn_fast", "k_grid_query", "k_kdw")


def short_name(kname: str) -> str:
    base = kname.split("(")[0].replace("void ", "").replace("<unnamed>::", "").strip()
    return base.split("<")[0].split("::")[-1]


if os.path.exists(launches):
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
    # one timed step = the launches from one k_qbin_count to the next; report one device-resident 1M-query step
    starts = [i for i, (nm, _) in enumerate(seq) if nm.endswith(STEP_FIRST)]
    groups = [seq[a:b] for a, b in zip(starts, starts[1:])]
    groups = [g for g in groups if not any("k_split" in nm or "k_level" in nm for nm, _ in g)]  # not a map build
    fast = [g for g in groups if any("k_knn_fast<0>" in nm for nm, _ in g)]  # the C2 pm_knn steps, not gathers
    groups = fast or groups
    # (the longest one: bench.py's later end-to-end steps run the same kernels on 131 072-query chunks)
    step = max(groups, key=lambda g: sum(t for _, t in g)) if groups else []
    with open(os.path.join(out, f"{tag}_step_launches.csv"), "w") as f:
        f.write("# launches of ONE pm_knn step (k_qbin_count .. last query kernel; the trailing fill is bench.py's untimed L2 flush), "
                "in launch order; same ncu pass as the full list\n")
        f.write("kernel,us,share_of_step\n")
        st = sum(t for _, t in step) or 1.0
        for nm, t in step:
            f.write(f"\"{nm[:80]}\",{t:.1f},{t / st:.4f}\n")
    with open(os.path.join(out, f"{tag}_launches.csv"), "w") as f:
        f.write("# ncu --metrics gpu__time_duration.sum --clock-control none (cold-cache, serialised launches: compare shares)\n")
        f.write("kernel,launches,total_us,share\n")
        for k, (c, t) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
            f.write(f"\"{k}\",{c},{t:.1f},{t / tot:.4f}\n")
    print(open(os.path.join(out, f"{tag}_step_launches.csv")).read())

want = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "dram__throughput.avg.pct_of_peak_sustained_elapsed",
        "lts__t_bytes.sum", "lts__t_sector_hit_rate.pct", "l1tex__t_sector_hit_rate.pct", "l1tex__t_bytes.sum",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "launch__registers_per_thread", "launch__grid_size", "launch__block_size",
        "smsp__inst_executed.sum", "sm__inst_executed.avg.per_cycle_active", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
        "smsp__thread_inst_executed_per_inst_executed.ratio", "sm__cycles_elapsed.avg", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
        "lts__throughput.avg.pct_of_peak_sustained_elapsed", "l1tex__throughput.avg.pct_of_peak_sustained_elapsed"]


def to_bytes(v, u):
    x = float(v.replace(",", ""))
    return x * {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}[u]


tr = os.path.join(out, "roofline_traffic.json")
j = json.load(open(tr)) if os.path.exists(tr) else {}
for rep in reps:
    raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    R = list(csv.reader(raw.splitlines()))
    if len(R) < 3:
        print("empty report", rep)
        continue
    Hh, U = R[0], R[1]
    per = collections.OrderedDict()  # short kernel name -> list of (time us, dram bytes) per captured launch
    first = {}
    for V in R[2:]:
        if len(V) != len(Hh):
            continue
        kname = V[Hh.index("Kernel Name")] if "Kernel Name" in Hh else "kernel"
        short = short_name(kname)
        rd, wr = Hh.index("dram__bytes_read.sum"), Hh.index("dram__bytes_write.sum")
        ti = Hh.index("gpu__time_duration.sum")
        per.setdefault(short, []).append((float(V[ti].replace(",", "")) * {"ns": 1e-3, "us": 1, "ms": 1e3}.get(U[ti], 1),
                                          to_bytes(V[rd], U[rd]) + to_bytes(V[wr], U[wr])))
        if short not in first:
            first[short] = (kname, V)
    for short, (kname, V) in first.items():
        with open(os.path.join(out, f"{tag}_{short}_ncu.txt"), "w") as f:
            f.write(f"# ncu --set full --clock-control none, first captured launch of {kname}\n")
            f.write(f"# launches_captured {len(per[short])}  sum_time_us {sum(t for t, _ in per[short]):.1f}  "
                    f"sum_dram_bytes {int(sum(b for _, b in per[short]))}\n")
            for w in want:
                if w in Hh:
                    i = Hh.index(w)
                    f.write(f"{w:70s} {V[i]:>18s} {U[i]}\n")
            for i, hname in enumerate(Hh):
                if hname.startswith("smsp__average_warps_issue_stalled") and hname.endswith("per_issue_active.ratio"):
                    try:
                        if float(V[i].replace(",", "")) >= 0.15:
                            f.write(f"{hname:70s} {V[i]:>18s} {U[i]}\n")
                    except ValueError:
                        pass
        print(open(os.path.join(out, f"{tag}_{short}_ncu.txt")).read())
    if "k_knn_fast" in per:
        # one step: first captured launch of each of the step's kernels
        j["knn_step_k64_dram_bytes_per_launch"] = int(sum(per[k][0][1] for k in STEP_KERNELS if k in per))
        j["knn_step_kernels"] = {k: int(per[k][0][1]) for k in STEP_KERNELS if k in per}
        j["knn_step_source"] = f"profiles/{tag}_*_ncu.txt ({os.path.basename(rep)})"
    if "k_walk" in per:
        j["k_walk_dram_bytes_per_launch"] = int(per["k_walk"][-1][1])
        j["k_walk_source"] = f"profiles/{tag}_k_walk_ncu.txt (last captured launch: the 4 M-path walk of bench.py's stage extras)"
    if "k_level" in per:
        names = [k for k in per if k in ("k_extract", "k_split", "k_level", "k_finish_subtrees", "k_emit_nodes", "k_make_soa", "k_gather44",
                                         "k_cell_ids", "k_grid_fill", "k_cells_transpose", "k_root", "k_sample_queries", "k_kdw") or k.startswith("DeviceRadixSort")]
        j["map_build_10M_dram_bytes_per_launch"] = int(sum(b for k in names for _, b in per[k]))
        j["map_build_10M_kernels"] = {k: [len(per[k]), int(sum(b for _, b in per[k]))] for k in names}
        j["map_build_source"] = f"profiles/{tag}_k_level_ncu.txt ({os.path.basename(rep)}: every kernel of one pm_map_build)"
json.dump(j, open(tr, "w"), indent=1)
print(json.dumps(j, indent=1))
