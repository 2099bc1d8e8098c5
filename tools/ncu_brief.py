"""One line of headline metrics + top stall reasons per captured launch of an .ncu-rep.

    python tools/ncu_brief.py gpurun_out/prof.ncu-rep [kernel regex]
"""
import csv
import subprocess
import sys

rep = sys.argv[1]
cmd = ["ncu", "-i", rep, "--page", "raw", "--csv"]
if len(sys.argv) > 2:
    cmd += ["--kernel-name", "regex:" + sys.argv[2]]
rows = list(csv.reader(subprocess.run(cmd, capture_output=True, text=True).stdout.splitlines()))
h = rows[0]
col = {c: i for i, c in enumerate(h)}
for v in rows[2:]:
    g = lambda k: v[col[k]] if k in col else "?"    # noqa: E731
    print(f"{g('Kernel Name')[:48]:48s} grid {g('launch__grid_size'):>6s} x{g('launch__block_size'):>4s} regs {g('launch__registers_per_thread'):>3s} "
          f"{float(g('gpu__time_duration.sum')):8.1f} us  inst {float(g('smsp__inst_executed.sum')) / 1e6:7.1f}M  "
          f"issue {float(g('smsp__issue_active.avg.pct_of_peak_sustained_active')):4.1f}%  warps {float(g('sm__warps_active.avg.pct_of_peak_sustained_active')):4.1f}%  "
          f"dram {float(g('gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed')):4.1f}%  waves {g('launch__waves_per_multiprocessor')}")
    st = [(c.split("stalled_")[1], float(v[i].replace(",", ""))) for c, i in col.items()
          if "pcsamp_warps_issue_stalled" in c and "not_issued" not in c and v[i] not in ("", "n/a")]
    tot = sum(x for _, x in st) or 1
    print("    " + "  ".join(f"{k} {100 * x / tot:.0f}%" for k, x in sorted(st, key=lambda t: -t[1])[:7]))
