"""Per-CUDA-source-line instruction counts and stall samples of one kernel in an .ncu-rep
(needs --import-source on at capture time and -lineinfo at compile time).

    python tools/ncu_lines.py gpurun_out/prof.ncu-rep <kernel regex> [top N] [launch index]
"""
import csv
import subprocess
import sys

rep, rx = sys.argv[1], sys.argv[2]
top = int(sys.argv[3]) if len(sys.argv) > 3 else 40
skip = sys.argv[4] if len(sys.argv) > 4 else "0"
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass", "--kernel-name",
                      "regex:" + rx, "--launch-skip", skip, "--launch-count", "1"], capture_output=True, text=True).stdout
data, hdr, fname = [], None, ""
for r in csv.reader(out.splitlines()):
    if not r:
        continue
    if r[0] == "File Path":
        fname = r[1].rsplit("/", 1)[-1]
    elif r[0] == "Line No":
        hdr = r
        ci = hdr.index("Instructions Executed"); cs = hdr.index("# Samples")
    elif hdr and r[0].isdigit() and len(r) > ci:
        try:
            data.append((int(r[ci] or 0), int(r[cs] or 0), fname + ":" + r[0], r[1].strip()[:105]))
        except ValueError:
            pass
tot = sum(d[0] for d in data) or 1
tots = sum(d[1] for d in data) or 1
print(f"total warp instructions {tot}, samples {tots}")
for d in sorted(data, key=lambda d: -d[1])[:top]:
    print(f"{100 * d[0] / tot:5.1f}% inst {100 * d[1] / tots:5.1f}% smpl  {d[2]:>20s}  {d[3]}")
