"""Group an `ncu --metrics gpu__time_duration.sum --csv` launch list by kernel.

    python tools/summarise_launches.py gpurun_out/launches.csv > profiles/rNN_launches.md

ncu serialises launches and runs them cold-cache, so only the SHARES are comparable with the
CUDA-event stage timings bench.py prints (roofline.stages[*].share)."""
from __future__ import annotations

import collections
import csv
import re
import sys


def main(path):
    with open(path) as f:
        lines = [ln for ln in f if not ln.startswith("==")]
    agg = collections.OrderedDict()
    for row in csv.DictReader(lines):
        if row.get("Metric Name") != "gpu__time_duration.sum":
            continue
        name = re.sub(r"\(.*", "", row["Kernel Name"]).replace("uwie::<unnamed>::", "").replace("void ", "")
        v = float(row["Metric Value"].replace(",", ""))
        if row["Metric Unit"] in ("ns", "nsecond"):
            v /= 1e3
        elif row["Metric Unit"] in ("ms", "msecond"):
            v *= 1e3
        a = agg.setdefault(name, [0, 0.0, row["Grid Size"], row["Block Size"]])
        a[0] += 1
        a[1] += v
    tot = sum(a[1] for a in agg.values()) or 1.0
    print(f"source: {path}\n")
    print("| kernel | launches | total us | share | grid (first launch) | block |")
    print("|---|---:|---:|---:|---|---|")
    for k, (n, t, g, b) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        print(f"| `{k[:70]}` | {n} | {t:.1f} | {100 * t / tot:.1f}% | {g} | {b} |")
    print(f"\ntotal {tot / 1e3:.2f} ms over {sum(a[0] for a in agg.values())} launches")


if __name__ == "__main__":
    main(sys.argv[1])
