"""Opcode mix (executed warp instructions and stall samples) of one kernel in an .ncu-rep.

    python tools/ncu_opmix.py gpurun_out/prof.ncu-rep <kernel regex> [top N]
"""
import collections
import csv
import subprocess
import sys

rep, rx = sys.argv[1], sys.argv[2]
top = int(sys.argv[3]) if len(sys.argv) > 3 else 30
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "sass", "--kernel-name",
                      "regex:" + rx, "--launch-count", "1"], capture_output=True, text=True).stdout
op, smp, ci = collections.Counter(), collections.Counter(), None
for r in csv.reader(out.splitlines()):
    if r and r[0] == "Address":
        ci, cs = r.index("Instructions Executed"), r.index("# Samples")
        continue
    if ci is None or len(r) <= ci or not r[0].startswith("0x"):
        continue
    s = r[1].strip()
    if s.startswith("@"):
        s = s.split(None, 1)[1]
    o = s.split()[0].split(".")[0]
    op[o] += int(r[ci] or 0)
    smp[o] += int(r[cs] or 0)
tot, ts = sum(op.values()) or 1, sum(smp.values()) or 1
print(f"total warp instructions {tot}, samples {ts}")
for o, n in op.most_common(top):
    print(f"{o:10s} {100 * n / tot:5.1f}% inst {100 * smp[o] / ts:5.1f}% smpl")
