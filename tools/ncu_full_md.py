"""Markdown summary of an `ncu --set full` report: one section per distinct (kernel, grid), first launch of each.

    python tools/ncu_full_md.py gpurun_out/prof.ncu-rep "title line" > profiles/rNN_ncu_full.md
"""
import csv
import subprocess
import sys

rep, title = sys.argv[1], sys.argv[2]
raw = open(rep).read() if rep.endswith(".csv") else subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(raw.splitlines()))
h, units = rows[0], rows[1]
col = {c: i for i, c in enumerate(h)}
keys = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
        "smsp__issue_active.avg.pct_of_peak_sustained_active", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "launch__registers_per_thread", "launch__shared_mem_per_block_static", "launch__shared_mem_per_block_dynamic",
        "launch__waves_per_multiprocessor", "smsp__inst_executed.sum"]
print("# " + title + "\n")
seen = set()
for v in rows[2:]:
    name = v[col["Kernel Name"]]
    grid = v[col["launch__grid_size"]]
    if (name, grid) in seen:
        continue
    seen.add((name, grid))
    print(f"## {name}  grid {v[col['Grid Size']] if 'Grid Size' in col else grid} block {v[col['Block Size']] if 'Block Size' in col else v[col['launch__block_size']]}")
    for k in keys:
        if k in col:
            print(f"- {k}: {v[col[k]]} {units[col[k]]}")
    st = [(c.split("stalled_")[1], float(v[i].replace(",", ""))) for c, i in col.items()
          if "pcsamp_warps_issue_stalled" in c and "not_issued" not in c and v[i] not in ("", "n/a")]
    tot = sum(x for _, x in st) or 1
    print("- stall samples: " + ", ".join(f"{k} {100 * x / tot:.0f} %" for k, x in sorted(st, key=lambda t: -t[1])[:6]))
    print()
