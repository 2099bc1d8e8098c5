"""Bench of the enhance + score hot path (BASELINE.json: 1080p images/sec, enhance + score).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl native|reference]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N ... bench.py --gpus N ...

One step = one pass of the six-strategy library (six_stadigy.py:392-447 loop body + scoring of every
output, quality_assessment.py:215-286) over a batch of 256 synthetic 1920x1080 frames PER GPU
(BASELINE.json configs[1]); frames are processed in chunks of --chunk frames so the working set of a
chunk stays a few GB (16 by default: fewer, fuller launches; the host entry point uses 8).  `value` times the device path with frames already in HBM; `e2e`
times the C-ABI host entry point (uwie_enhance_score_host) with host buffers: H2D of the frames,
kernels, D2H of all six enhanced frames + score records.  Ranks share nothing: images are
partitioned by index, and the only exchange is the NCCL all-gather of the per-image score records.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

H, W = 1080, 1920
NPIX = H * W
METRIC = "1080p images/sec (enhance+score)"
UNIT = "images/s"


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="native", choices=("native", "reference"))
    ap.add_argument("--batch", type=int, default=256, help="frames per GPU per step")
    ap.add_argument("--chunk", type=int, default=16, help="frames per engine call (device path)")
    ap.add_argument("--e2e-chunk", type=int, default=8, help="frames per engine call inside the host entry point")
    ap.add_argument("--distinct", type=int, default=24, help="distinct seeded frames (tiled to --batch)")
    ap.add_argument("--library", default="six", choices=("six", "es"))
    ap.add_argument("--lanes", type=int, default=4,
                    help="engines (own workspace + stream each) that take the chunks of a step in turn")
    ap.add_argument("--e2e-frames", type=int, default=256, help="frames per e2e step (host buffers)")
    ap.add_argument("--cpu-frames", type=int, default=0, help="cpu_baseline sample size (0 = one per worker)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-numa-bind", action="store_true", help="N > 1: do not pin ranks to their GPU's NUMA node")
    ap.add_argument("--entry-points", action="store_true",
                    help="instead of the headline line: device timing of the other entry points (configs[2], "
                         "configs[3] pieces, features79, quality_scores) with the CPU restatement beside each")
    ap.add_argument("--out", default="", help="--entry-points: also write the JSON lines to this file")
    return ap.parse_args()


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        try:
            return float(json.load(open(p))["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
        except Exception:
            pass
    return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


# ------------------------------------------------------------------------------------------
# clocks sampler (nvidia-smi during the timed region)
# ------------------------------------------------------------------------------------------
class ClockSampler:
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index: int):
        self.index, self.samples, self._stop, self._t = index, [], threading.Event(), None

    def _loop(self):
        while not self._stop.is_set():
            try:
                out = subprocess.run(["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.Q}",
                                      "--format=csv,noheader,nounits"], capture_output=True, text=True,
                                     timeout=5).stdout.strip()
                if out:
                    self.samples.append([x.strip() for x in out.split(",")])
            except Exception:
                pass
            self._stop.wait(0.2)

    def __enter__(self):
        self._t = threading.Thread(target=self._loop, daemon=True)
        self._t.start()
        return self

    def __exit__(self, *a):
        self._stop.set()
        self._t.join(timeout=6)

    def summary(self):
        sm, mx, reasons = [], [], set()
        names = ("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap")
        for s in self.samples:
            try:
                sm.append(float(s[0])); mx.append(float(s[1]))
            except Exception:
                continue
            for n, v in zip(names, s[2:6]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        if not sm:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
        return {"sm_mhz": float(np.median(sm)), "sm_max_mhz": float(max(mx)), "reasons": sorted(reasons),
                "samples": len(sm)}


# ------------------------------------------------------------------------------------------
# CPU arm: the oracle restatement of the reference's NumPy/OpenCV path on the host cores
# (the reference itself is Python under /root/reference, which does not exist on the GPU box;
#  oracle/restatement.py is pinned bit-for-bit to it by tests/test_oracle_golden.py -> kind "port")
# ------------------------------------------------------------------------------------------
def _cpu_worker(job):
    first_seed, count, library = job
    import cv2
    cv2.setNumThreads(1)
    from oracle import restatement as o
    from underwater_image_enhancement_b200.synth import synth_image
    fn = o.p6_frame if library == "six" else o.p5_frame
    frames = [synth_image(first_seed + i, H, W) for i in range(count)]
    t0 = time.perf_counter()
    for f in frames:
        fn(f)
    return time.perf_counter() - t0


def host_cores():
    try:
        n = len(os.sched_getaffinity(0))
    except Exception:
        n = os.cpu_count() or 1
    return max(1, min(n, 64))


def cpu_throughput(library: str, workers: int, frames_per_worker: int, first_seed: int = 0):
    """images/s of the CPU path with `workers` processes, one OpenCV thread each."""
    import multiprocessing as mp
    ctx = mp.get_context("spawn")
    jobs = [(first_seed + i * frames_per_worker, frames_per_worker, library) for i in range(workers)]
    with ctx.Pool(workers) as pool:
        busy = pool.map(_cpu_worker, jobs)
    return workers * frames_per_worker / max(busy), max(busy)


def run_reference_arm(args):
    """bench.py --impl reference: the reference's CPU implementation of the path (oracle port) with
    all host cores; each step is a bounded sample (one frame per worker process)."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    workers = host_cores()
    per = max(1, args.cpu_frames // workers) if args.cpu_frames else 1
    for i in range(min(args.warmup, 1)):           # a CPU path has no caches worth 3 warm-ups of 6 s/frame
        cpu_throughput(args.library, workers, per, first_seed=1000)
    vals, secs = [], 0.0
    for k in range(args.steps):
        v, s = cpu_throughput(args.library, workers, per, first_seed=k * workers * per)
        vals.append(v); secs += s
    value = workers * per * args.steps / secs
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * secs / max(args.steps, 1),
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "u8+f32/f64",
        "data": "synthetic",
        "config": {"workload": f"{args.library}-strategy enhance+score, 1920x1080 synthetic frames (configs[1])",
                   "frames_per_step": workers * per, "library": args.library},
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": workers, "kind": "port",
                         "sample": f"{workers * per} frames per step x {args.steps} steps, one process per core, "
                                   "cv2.setNumThreads(1); oracle/restatement.py (pinned to the reference by golden fixtures)"},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)
    return 0


# ------------------------------------------------------------------------------------------
# native arm
# ------------------------------------------------------------------------------------------
def stage_alg_bytes(library: str):
    """Algorithmic (compulsory in + out) bytes PER FRAME of each stage of the kernel sequence, in
    units of N = H*W pixels (DESIGN.md 'Kernels and their rooflines').  six: S=6 strategies, 3 of
    them dehaze (float32 recovered planes); es: S=5, 3 dehaze (float64 recovered planes)."""
    S, nd, fsz = (6, 3, 4) if library == "six" else (5, 3, 8)
    n_clahe_dehaze = 2 if library == "six" else 0
    n_clahe_frame = 3 if library == "six" else 1
    rec = 3 * fsz                                             # bytes per pixel of one recovered frame
    return {
        "frame_hist": 3,                                      # read the frame once
        "quadtree_canny": 4,                                  # frame quadrants: 3N * (1 + 1/4 + 1/16 + ...)
        "guided_recover": 3 + nd * rec,                       # frame in, recovered planes out
        "percentile_select": nd * rec,                        # one read of the data being ranked
        "clahe_tile_luts": n_clahe_dehaze * rec + n_clahe_frame * 3,
        "clahe_out_hist": 3,
        "final_apply": nd * rec + (S - nd) * 3 + S * 4,       # stage inputs in, S x (RGB + gray) out
        "score_canny": S * 1,                                 # gray planes in, counts out
    }


def bind_to_gpu_numa(local):
    """Pin this rank (and the pinned host buffers it allocates afterwards) to the CPUs of its GPU's NUMA
    node, so that the host<->device copies of the e2e arm do not cross the socket interconnect when several
    ranks share the box.  Best effort: returns a description, or None when the topology cannot be read."""
    try:
        import pynvml
        pynvml.nvmlInit()
        vis = os.environ.get("CUDA_VISIBLE_DEVICES")
        if vis:
            tok = vis.split(",")[local].strip()
            h = pynvml.nvmlDeviceGetHandleByUUID(tok) if tok.startswith("GPU-") else pynvml.nvmlDeviceGetHandleByIndex(int(tok))
        else:
            h = pynvml.nvmlDeviceGetHandleByIndex(local)
        bus = pynvml.nvmlDeviceGetPciInfo(h).busId
        bus = bus.decode() if isinstance(bus, bytes) else bus
        path = "/sys/bus/pci/devices/" + bus.lower()[-12:]
        node = int(open(path + "/numa_node").read())
        cpus = set()
        for part in open(path + "/local_cpulist").read().strip().split(","):
            lo, _, hi = part.partition("-")
            cpus.update(range(int(lo), int(hi or lo) + 1))
        cpus &= os.sched_getaffinity(0)
        if not cpus:
            return None
        os.sched_setaffinity(0, cpus)
        return {"node": node, "cpus": len(cpus)}
    except Exception:                                                              # noqa: BLE001
        return None


def run_native(args):
    import torch
    import torch.distributed as dist

    from underwater_image_enhancement_b200 import _lib, shard
    from underwater_image_enhancement_b200.engine import Engine
    from underwater_image_enhancement_b200.synth import synth_image

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise RuntimeError("bench.py (native arm) needs a CUDA device: there is no CPU fallback")
    numa = bind_to_gpu_numa(local) if (world > 1 and not args.no_numa_bind) else None
    print(f"[bench] rank {rank} numa binding: {numa}", file=sys.stderr, flush=True)
    torch.cuda.set_device(local)
    dev = torch.device(f"cuda:{local}")
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    _lib.load()                                                # fails loudly if the .so is missing

    B, chunk = args.batch, min(args.chunk, args.batch)
    S = 6 if args.library == "six" else 5
    distinct = min(args.distinct, B)
    host = np.stack([synth_image(rank * distinct + i, H, W) for i in range(distinct)])
    idx = torch.arange(B, device=dev) % distinct
    d_frames = torch.from_numpy(host).to(dev)[idx].contiguous()          # [B,H,W,3] resident in HBM
    out_all = torch.empty((B, S, H, W, 3), dtype=torch.uint8, device=dev)
    scores = torch.empty((B, S, 9), dtype=torch.float64, device=dev)
    best = torch.empty((B,), dtype=torch.int32, device=dev)
    types = torch.empty((B,), dtype=torch.int32, device=dev)
    lanes = max(1, args.lanes)
    engs = [Engine(args.library, H, W, max_batch=chunk, device=str(dev)) for _ in range(lanes)]
    eng = engs[0]
    side = [torch.cuda.Stream(device=dev) for _ in range(lanes)] if lanes > 1 else []

    launches = [0]

    def run_chunks(lane_engs, lane_streams):
        main = torch.cuda.current_stream(dev)
        for st in lane_streams:
            st.wait_stream(main)
        for n, c0 in enumerate(range(0, B, chunk)):
            c1 = min(B, c0 + chunk)
            e = lane_engs[n % len(lane_engs)]
            if lane_streams:
                with torch.cuda.stream(lane_streams[n % len(lane_streams)]):
                    e.run(d_frames[c0:c1], want_all=True, out_all=out_all[c0:c1], scores=scores[c0:c1],
                          best=best[c0:c1], types=types[c0:c1])
            else:
                e.run(d_frames[c0:c1], want_all=True, out_all=out_all[c0:c1], scores=scores[c0:c1],
                      best=best[c0:c1], types=types[c0:c1])
            launches[0] += e.last_launches
        for st in lane_streams:
            main.wait_stream(st)

    def step():
        run_chunks(engs, side)
        if world > 1:      # the path's only exchange: per-image score records to every rank (rank 0 reads them)
            shard.gather_records(shard.pack_records(scores, best, types, B))

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize(dev)

    for _ in range(max(args.warmup, 0)):
        step()
    eng.check()
    barrier()
    launches[0] = 0
    if lanes == 1:
        eng.profile(True)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    with ClockSampler(local) as clk:
        barrier()
        e0.record()
        for _ in range(args.steps):
            step()
        e1.record()
        barrier()
    ms = e0.elapsed_time(e1)
    stage_note = "CUDA events around every stage, inside the timed region"
    if lanes > 1:
        # with several lanes the stages of different chunks overlap on the device, so their event
        # brackets are not additive; the per-stage table comes from one more pass on a single lane
        eng.profile(True)
        timed_launches = launches[0]
        for _ in range(args.steps):
            run_chunks([eng], [])
        launches[0] = timed_launches
        stage_note = f"second pass of the same {args.steps} steps on ONE lane (stages of different lanes overlap in the timed region)"
    stages = eng.profile_read()
    eng.profile(False)
    for e in engs:
        e.check()
    t = torch.tensor([ms], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms_max = float(t.item())
    value = world * B * args.steps / (ms_max / 1e3)

    # ---- e2e: C-ABI host entry point, pinned host buffers, H2D + kernels + D2H inside the timed region
    ne = min(args.e2e_frames, B)
    h_frames = torch.empty((ne, H, W, 3), dtype=torch.uint8, pin_memory=True)
    h_frames.copy_(torch.from_numpy(host)[torch.arange(ne) % distinct])
    h_out = torch.empty((ne, S, H, W, 3), dtype=torch.uint8, pin_memory=True)
    # the host entry point pipelines copy-in / kernels / copy-out chunk by chunk: short chunks keep the
    # exposed first copy-in and last copy-out small, so it gets its own engine
    e2e_chunk = max(1, min(args.e2e_chunk, ne))
    heng = eng if e2e_chunk == chunk else Engine(args.library, H, W, max_batch=e2e_chunk, device=str(dev))
    for _ in range(2):
        res = heng.run_host(h_frames.numpy(), want_all=True, out_all=h_out.numpy())
    barrier()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        res = heng.run_host(h_frames.numpy(), want_all=True, out_all=h_out.numpy())
    torch.cuda.synchronize(dev)
    e2e_s = time.perf_counter() - t0
    t = torch.tensor([e2e_s], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    e2e_value = world * ne * args.steps / float(t.item())
    h2d = ne * NPIX * 3
    d2h = ne * S * NPIX * 3 + ne * (S * 9 * 8 + 4 + 4 + 12)
    # the host path must agree with the device path on the same frames
    same = np.array_equal(res["scores"], scores[:ne].cpu().numpy()) and \
        np.array_equal(h_out.numpy()[0], out_all[0].cpu().numpy())

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return 0

    peak, peak_src = peaks()
    alg = stage_alg_bytes(args.library)
    total_stage_ms = sum(m for _, m, _, _ in stages) or 1.0
    table = []
    for name, m, ker, runs in stages:
        per_frame = alg.get(name)
        frames_done = B * args.steps
        gbs = (per_frame * NPIX * frames_done / (m / 1e3) / 1e9) if (per_frame and m > 0) else None
        table.append({"stage": name, "ms_per_step": m / args.steps, "share": m / total_stage_ms,
                      "kernels_per_call": ker // max(runs, 1), "alg_bytes_per_frame": (per_frame or 0) * NPIX,
                      "achieved_gbs": gbs})
    dom = max((r for r in table if r["achieved_gbs"]), key=lambda r: r["ms_per_step"])
    traffic = None
    tp = os.path.join(ROOT, "profiles", "roofline_traffic.json")
    if os.path.exists(tp):
        try:
            traffic = json.load(open(tp)).get(dom["stage"])
        except Exception:
            traffic = None
    calls_per_step = (B + chunk - 1) // chunk
    roofline = {
        "bound": "hbm", "kernel": dom["stage"], "achieved": dom["achieved_gbs"], "peak": peak, "unit": "GB/s",
        "frac": dom["achieved_gbs"] / peak, "traffic": traffic, "peak_source": peak_src,
        "alg_bytes_per_launch": dom["alg_bytes_per_frame"] * chunk,
        "avg_launch_ms": dom["ms_per_step"] / calls_per_step,
        "share_of_step": dom["share"],
        "pipeline": {"alg_bytes_per_frame": (3 + 3 * S) * NPIX,
                     "achieved": (3 + 3 * S) * NPIX * value / world / 1e9,
                     "frac": (3 + 3 * S) * NPIX * value / world / 1e9 / peak},
        "stages": table, "stage_timing": stage_note,
    }
    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": ms_max / args.steps, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "u8+f32" if args.library == "six" else "u8+f64",
        "data": "synthetic",
        "config": {"workload": f"{args.library}-strategy enhance+score, batch of {B} 1920x1080 synthetic frames per GPU "
                               "(BASELINE.json configs[1])",
                   "library": args.library, "frames_per_gpu_per_step": B, "chunk": chunk,
                   "distinct_frames": distinct, "strategies": S, "lanes": lanes,
                   "l2": f"inputs {B * NPIX * 3 / 1e6:.0f} MB + outputs {B * S * NPIX * 3 / 1e6:.0f} MB per step exceed the 126 MB L2; no flush needed",
                   "parallelism": f"image-index sharding x{world}, NCCL all-gather of score records",
                   "numa_bind": numa},
        "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                "frames_per_step": ne, "chunk": e2e_chunk, "api": "uwie_enhance_score_host (C ABI, pinned host buffers)",
                "matches_device_path": bool(same)},
        "gpu_launches": launches[0],
        "clocks": clk.summary(),
        "roofline": roofline,
    }
    if world == 1 and not args.no_cpu_baseline:
        workers = host_cores()
        per = max(1, args.cpu_frames // workers) if args.cpu_frames else 1
        v, busy = cpu_throughput(args.library, workers, per)
        line["cpu_baseline"] = {"value": v, "unit": UNIT, "cores": workers, "kind": "port",
                                "sample": f"{workers * per} frames of the same workload ({busy:.1f} s), one process per core, "
                                          "oracle/restatement.py (NumPy/OpenCV restatement pinned to the reference)"}
    print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()
    return 0


def run_entry_points(args):
    """tools/bench_entry_points.py with this file's cpu_baseline leg supplying the CPU side (oracle port,
    one frame per entry on one host core)."""
    sys.path.insert(0, os.path.join(ROOT, "tools"))
    import bench_entry_points as bep
    cpu = None
    if not args.no_cpu_baseline:
        from oracle import restatement as o

        def es_scored(img, name, params):
            o.qa_total(o.es_strategy(img, name, params))

        cpu = {"es_strategy_scored": es_scored, "vgg_enhance": o.vgg_enhance, "vgg_features": o.vgg_features,
               "vgg_preprocess": o.vgg_preprocess, "features79": o.fe_all, "quality_scores": o.qa_total}
    bep.main(cpu=cpu, out_path=args.out)


def main():
    args = parse_args()
    if args.entry_points:
        run_entry_points(args)
        return
    if args.impl == "reference":
        return run_reference_arm(args)
    return run_native(args)


if __name__ == "__main__":
    sys.exit(main())
