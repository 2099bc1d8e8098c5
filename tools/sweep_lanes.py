"""bench.py device value over lanes x chunk (GPU box): python tools/sweep_lanes.py > gpurun_out/sweep.txt"""
import json
import subprocess
import sys

for lanes in (2, 3, 4, 6, 8):
    for chunk in (8, 16, 32):
        r = subprocess.run([sys.executable, "bench.py", "--steps", "2", "--warmup", "2", "--no-cpu-baseline", "--probe-mb", "0",
                            "--no-p5", "--lanes", str(lanes), "--chunk", str(chunk), "--e2e-frames", "16"],
                           capture_output=True, text=True)
        for ln in r.stdout.splitlines():
            if ln.startswith("{"):
                d = json.loads(ln)
                print(f"lanes {lanes} chunk {chunk}: {d['value']:.1f} frames/s ({d['ms_per_step']:.1f} ms)", flush=True)
