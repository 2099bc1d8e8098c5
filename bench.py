"""Bench of the enhance + score hot path (BASELINE.json: 1080p images/sec, enhance + score).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl native|reference]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N ... bench.py --gpus N ...

One step = one pass of the six-strategy library (six_stadigy.py:392-447 loop body + scoring of every
output, quality_assessment.py:215-286) over a batch of 256 synthetic 1920x1080 frames PER GPU
(BASELINE.json configs[1], "P6": 21 N algorithmic bytes per frame); frames are processed in chunks of
--chunk frames so the working set of a chunk stays a few GB.  `value` times the device path with frames
already in HBM; `e2e` times the C-ABI host entry point (uwie_enhance_score_host) with host buffers: H2D of
the frames, kernels, D2H of all six enhanced frames + score records.  The same JSON line carries a `p5`
record: the five-strategy library of enhancement_strategies.py under main.py's best-only contract
(main.py:122-155: frame in, selected frame + scores out, 6 N bytes), device-timed and end to end.
Ranks share nothing: images are partitioned by index, and the only exchange is the NCCL all-gather of the
per-image score records, whose content rank 0 checks (`gather_checked`).
"""
from __future__ import annotations

import argparse
import ctypes
import glob
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


def workload_name(library: str, batch: int) -> str:
    """config.workload: the SAME string for the native and the reference arm."""
    return (f"{library}-strategy enhance+score, batch of {batch} 1920x1080 synthetic frames per GPU "
            "(BASELINE.json configs[1])")


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
    ap.add_argument("--no-p5", action="store_true", help="skip the p5 record (ES library, best-only contract)")
    ap.add_argument("--no-numa-bind", action="store_true", help="N > 1: do not pin ranks to their GPU's NUMA node")
    ap.add_argument("--probe-mb", type=int, default=512, help="size of the per-rank PCIe probe copies (0 = skip)")
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
# CPU arm.  kind "reference": the UNMODIFIED reference modules from baseline/_ref (a verbatim, git-ignored
# copy of /root/reference made by __graft_entry__.build(); it travels to the GPU box with gpurun) driven by
# oracle/refload.py.  kind "port": oracle/restatement.py (pinned bit for bit to the reference by
# tests/test_oracle_golden.py) when that copy is absent.
# ------------------------------------------------------------------------------------------
def cpu_kind():
    from oracle import refload
    return "reference" if refload.reference_dir() else "port"


def _cpu_worker(job):
    first_seed, count, library, kind = job
    import cv2
    cv2.setNumThreads(1)
    from underwater_image_enhancement_b200.synth import synth_image
    if kind == "reference":
        from oracle import refload
        mods = refload.load_reference()
        fn = (lambda f: refload.p6_frame(mods, f)) if library == "six" else (lambda f: refload.p5_frame(mods, f))
    else:
        from oracle import restatement as o
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


def cpu_throughput(library: str, workers: int, frames_per_worker: int, first_seed: int = 0, kind: str = "port"):
    """images/s of the CPU path with `workers` processes, one OpenCV thread each."""
    import multiprocessing as mp
    ctx = mp.get_context("spawn")
    jobs = [(first_seed + i * frames_per_worker, frames_per_worker, library, kind) for i in range(workers)]
    with ctx.Pool(workers) as pool:
        busy = pool.map(_cpu_worker, jobs)
    return workers * frames_per_worker / max(busy), max(busy)


def cpu_sample_text(kind, workers, per, steps=None):
    what = ("the unmodified reference (baseline/_ref: six_stadigy.py / enhancement_strategies.py / "
            "quality_assessment.py, skimage.measure.shannon_entropy stand-in)") if kind == "reference" else \
           "oracle/restatement.py (NumPy/OpenCV restatement pinned bit for bit to the reference)"
    n = f"{workers * per} frames per step x {steps} steps" if steps else f"{workers * per} frames"
    return f"{n} of the same workload, one process per core, cv2.setNumThreads(1); {what}"


def run_reference_arm(args):
    """bench.py --impl reference: the reference's own CPU implementation of the path with all host cores;
    each step is a bounded sample (one frame per worker process) of the native arm's workload."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    kind = cpu_kind()
    workers = host_cores()
    per = max(1, args.cpu_frames // workers) if args.cpu_frames else 1
    for i in range(min(args.warmup, 1)):           # a CPU path has no caches worth 3 warm-ups of 6 s/frame
        cpu_throughput(args.library, workers, per, first_seed=1000, kind=kind)
    secs = 0.0
    for k in range(args.steps):
        _, s = cpu_throughput(args.library, workers, per, first_seed=k * workers * per, kind=kind)
        secs += s
    value = workers * per * args.steps / secs
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * secs / max(args.steps, 1),
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "u8+f32" if args.library == "six" else "u8+f64", "data": "synthetic",
        "config": {"workload": workload_name(args.library, args.batch), "library": args.library,
                   "frames_per_gpu_per_step": args.batch, "sample_frames_per_step": workers * per},
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": workers, "kind": kind,
                         "sample": cpu_sample_text(kind, workers, per, args.steps)},
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
        # CLAHE strategies: stage input in (recovered planes or frame bytes), Lab bytes out (the hand-off to the apply passes)
        "clahe_tile_luts": n_clahe_dehaze * rec + n_clahe_frame * 3 + (n_clahe_dehaze + n_clahe_frame) * 3,
        "clahe_out_hist": 3 + 3,                              # Lab bytes in, CLAHE output bytes out (one strategy)
        # CLAHE strategies continue from 3 N of Lab (or CLAHE) bytes, the others from their stage input; S x (RGB + gray) out
        "final_apply": (n_clahe_dehaze + n_clahe_frame) * 3 + (nd - n_clahe_dehaze) * rec
                       + (S - nd - n_clahe_frame) * 3 + S * 4,
        "score_canny": S * 1,                                 # gray planes in, counts out
    }


def survey_unit_bytes(library: str):
    """SURVEY.md §8(d) units that correspond to a stage: 'guided-filter dehaze (A given)' is 6 N per
    instance (u8 RGB in -> u8 RGB out) and a frame has three dehaze instances."""
    return {"guided_recover": 6 * 3}


# ---- NUMA: bind the rank (threads + the pinned buffers it allocates afterwards) to its GPU's node
def _parse_cpulist(text):
    cpus = set()
    for part in text.strip().split(","):
        if part:
            lo, _, hi = part.partition("-")
            cpus.update(range(int(lo), int(hi or lo) + 1))
    return cpus


def _gpu_bus_id(local):
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
    return bus.lower()[-12:], pynvml.nvmlDeviceGetIndex(h)


def _topo_numa(index):
    """'NUMA Affinity' column of `nvidia-smi topo -m` for GPU `index`, or None."""
    try:
        out = subprocess.run(["nvidia-smi", "topo", "-m"], capture_output=True, text=True, timeout=20).stdout
        lines = [ln for ln in out.splitlines() if ln.strip()]
        head = next(ln for ln in lines if "NUMA Affinity" in ln)
        cols = [c.strip() for c in head.split("\t") if c.strip()]
        col = cols.index("NUMA Affinity")
        for ln in lines:
            cells = [c.strip() for c in ln.split("\t") if c.strip()]
            if cells and cells[0] == f"GPU{index}":
                v = cells[col + 1] if len(cells) > col + 1 else ""       # first cell is the row label
                return int(v.split(",")[0].split("-")[0]) if v and v[0].isdigit() else None
    except Exception:                                                       # noqa: BLE001
        return None
    return None


def bind_to_gpu_numa(local, world):
    """Pin this rank's threads, and through the memory policy the pinned host buffers it allocates afterwards,
    to the NUMA node of its GPU so that the e2e copies do not cross the socket interconnect when several ranks
    share the box.  Node from /sys/bus/pci/devices/<bdf>/numa_node, else the NUMA Affinity column of
    `nvidia-smi topo -m`, else (several nodes, nothing known) ranks spread over the nodes in order.  Returns
    what was found and done; on a single-node (or NUMA-less VM) host there is nothing to bind."""
    info = {"host_nodes": 0, "node": None, "source": None, "cpus": None, "mempolicy": None}
    try:
        nodes = sorted(int(os.path.basename(p)[4:]) for p in glob.glob("/sys/devices/system/node/node[0-9]*"))
        info["host_nodes"] = len(nodes)
        if len(nodes) < 2:
            info["source"] = "single NUMA node visible: nothing to bind"
            return info
        bus, index = _gpu_bus_id(local)
        node = -1
        try:
            node = int(open(f"/sys/bus/pci/devices/{bus}/numa_node").read())
            info["source"] = "sysfs numa_node"
        except Exception:                                                   # noqa: BLE001
            pass
        if node < 0:
            t = _topo_numa(index)
            if t is not None:
                node, info["source"] = t, "nvidia-smi topo -m"
        if node < 0:
            node, info["source"] = nodes[(local * len(nodes)) // max(world, 1)], "ranks spread over nodes in order"
        info["node"] = node
        cpus = _parse_cpulist(open(f"/sys/devices/system/node/node{node}/cpulist").read()) & os.sched_getaffinity(0)
        if cpus:
            os.sched_setaffinity(0, cpus)
            info["cpus"] = len(cpus)
        # MPOL_PREFERRED on that node for every later allocation of this process (cudaHostAlloc included)
        libc = ctypes.CDLL(None, use_errno=True)
        mask = (ctypes.c_ulong * 16)()
        mask[node // 64] = 1 << (node % 64)
        rc = libc.syscall(238, 1, ctypes.byref(mask), 16 * 64)              # SYS_set_mempolicy, MPOL_PREFERRED
        info["mempolicy"] = "preferred" if rc == 0 else f"set_mempolicy errno {ctypes.get_errno()}"
    except Exception as e:                                                  # noqa: BLE001
        info["source"] = f"failed: {e}"
    return info


def pcie_probe(torch, dist, dev, world, mb):
    """Per-rank host<->device copy bandwidth with ALL ranks copying at the same time (pinned buffers, CUDA
    events): D2H alone, H2D alone, both directions at once.  Returns GB/s lists indexed by rank (rank 0)."""
    if mb <= 0:
        return None
    n = mb << 20
    hbuf = torch.empty(n, dtype=torch.uint8, pin_memory=True)
    hbuf2 = torch.empty(n, dtype=torch.uint8, pin_memory=True)
    dbuf = torch.empty(n, dtype=torch.uint8, device=dev)
    dbuf2 = torch.empty(n, dtype=torch.uint8, device=dev)
    s1, s2 = torch.cuda.Stream(device=dev), torch.cuda.Stream(device=dev)
    reps = 4

    def timed(fn):
        fn()
        torch.cuda.synchronize(dev)
        if world > 1:
            dist.barrier()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        for _ in range(reps):
            fn()
        s1.synchronize(); s2.synchronize()
        b.record()
        torch.cuda.synchronize(dev)
        return reps * n / (a.elapsed_time(b) / 1e3) / 1e9

    def d2h():
        with torch.cuda.stream(s1):
            hbuf.copy_(dbuf, non_blocking=True)

    def h2d():
        with torch.cuda.stream(s2):
            dbuf2.copy_(hbuf2, non_blocking=True)

    def both():
        d2h(); h2d()

    # events are recorded on the current stream: make it wait for the copy streams
    def wrap(fn):
        def g():
            cur = torch.cuda.current_stream(dev)
            s1.wait_stream(cur); s2.wait_stream(cur)
            fn()
            cur.wait_stream(s1); cur.wait_stream(s2)
        return g

    vals = torch.tensor([timed(wrap(d2h)), timed(wrap(h2d)), timed(wrap(both))], dtype=torch.float64, device=dev)
    if world > 1:
        allv = [torch.zeros_like(vals) for _ in range(world)]
        dist.all_gather(allv, vals)
    else:
        allv = [vals]
    m = torch.stack(allv).cpu().numpy()
    return {"mb_per_copy": mb, "ranks_copying_at_once": world,
            "d2h_gbs": [round(float(x), 2) for x in m[:, 0]], "h2d_gbs": [round(float(x), 2) for x in m[:, 1]],
            "bidir_each_way_gbs": [round(float(x), 2) for x in m[:, 2]],
            "aggregate_d2h_gbs": round(float(m[:, 0].sum()), 1)}


def load_traffic():
    """profiles/roofline_traffic.json: per-stage dram read+write bytes per engine call, written by
    tools/traffic_from_ncu.py from the `ncu --set full` capture named in its `_source` key."""
    tp = os.path.join(ROOT, "profiles", "roofline_traffic.json")
    try:
        return json.load(open(tp))
    except Exception:                                                       # noqa: BLE001
        return {}


def measure_library(torch, dist, args, library, contract, dev, rank, world, host):
    """Device-timed value, per-stage table and e2e of one library under one output contract
    ('all': every strategy output is a result, P6; 'best': the selected output only, P5)."""
    from underwater_image_enhancement_b200 import shard
    from underwater_image_enhancement_b200.engine import Engine

    B, chunk = args.batch, min(args.chunk, args.batch)
    S = 6 if library == "six" else 5
    distinct = host.shape[0]
    want_all = contract == "all"
    idx = torch.arange(B, device=dev) % distinct
    d_frames = torch.from_numpy(host).to(dev)[idx].contiguous()          # [B,H,W,3] resident in HBM
    out_all = torch.empty((B, S, H, W, 3), dtype=torch.uint8, device=dev) if want_all else None
    out_best = None if want_all else torch.empty((B, H, W, 3), dtype=torch.uint8, device=dev)
    scores = torch.empty((B, S, 9), dtype=torch.float64, device=dev)
    best = torch.empty((B,), dtype=torch.int32, device=dev)
    types = torch.empty((B,), dtype=torch.int32, device=dev)
    lanes = max(1, args.lanes)
    engs = [Engine(library, H, W, max_batch=chunk, device=str(dev)) for _ in range(lanes)]
    eng = engs[0]
    side = [torch.cuda.Stream(device=dev) for _ in range(lanes)] if lanes > 1 else []
    launches = [0]

    def one(e, c0, c1):
        e.run(d_frames[c0:c1], want_all=want_all, want_best=not want_all,
              out_all=None if out_all is None else out_all[c0:c1],
              out_best=None if out_best is None else out_best[c0:c1],
              scores=scores[c0:c1], best=best[c0:c1], types=types[c0:c1])
        launches[0] += e.last_launches

    def run_chunks(lane_engs, lane_streams):
        main = torch.cuda.current_stream(dev)
        for st in lane_streams:
            st.wait_stream(main)
        for n, c0 in enumerate(range(0, B, chunk)):
            c1 = min(B, c0 + chunk)
            e = lane_engs[n % len(lane_engs)]
            if lane_streams:
                with torch.cuda.stream(lane_streams[n % len(lane_streams)]):
                    one(e, c0, c1)
            else:
                one(e, c0, c1)
        for st in lane_streams:
            main.wait_stream(st)

    gathered = [None]

    def step():
        run_chunks(engs, side)
        if world > 1:      # the path's only exchange: per-image score records to every rank (rank 0 reads them)
            gathered[0] = shard.gather_records(shard.pack_records(scores, best, types, B))

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
    local = dev.index or 0
    with ClockSampler(local) as clk:
        barrier()
        e0.record()
        for _ in range(args.steps):
            step()
        e1.record()
        barrier()
    ms = e0.elapsed_time(e1)
    timed_launches = launches[0]
    stage_note = "CUDA events around every stage, inside the timed region"
    if lanes > 1:
        # with several lanes the stages of different chunks overlap on the device, so their event
        # brackets are not additive; the per-stage table comes from one more pass on a single lane
        eng.profile(True)
        for _ in range(args.steps):
            run_chunks([eng], [])
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

    # the all-gather's content: every rank's block of the gathered tensor must be that rank's own records
    gather_ok = None
    if world > 1:
        mine = shard.pack_records(scores, best, types, B)
        blocks = [torch.empty_like(mine) for _ in range(world)]
        dist.all_gather(blocks, mine)                                    # list form: an independent route
        g_scores, g_best, g_types = shard.unpack_records(gathered[0], S)
        ok = g_scores.shape[0] == world * B
        for r in range(world):
            ok = ok and torch.equal(gathered[0][r * B:(r + 1) * B], blocks[r])
        ok = ok and torch.equal(g_scores[rank * B:(rank + 1) * B], scores) and \
            torch.equal(g_best[rank * B:(rank + 1) * B], best) and torch.equal(g_types[rank * B:(rank + 1) * B], types)
        flag = torch.tensor([1 if ok else 0], device=dev)
        dist.all_reduce(flag, op=dist.ReduceOp.MIN)
        gather_ok = bool(flag.item())

    # ---- e2e: C-ABI host entry point, pinned host buffers, H2D + kernels + D2H inside the timed region
    ne = min(args.e2e_frames, B)
    h_frames = torch.empty((ne, H, W, 3), dtype=torch.uint8, pin_memory=True)
    h_frames.copy_(torch.from_numpy(host)[torch.arange(ne) % distinct])
    h_out = torch.empty((ne, S, H, W, 3) if want_all else (ne, H, W, 3), dtype=torch.uint8, pin_memory=True)
    # the host entry point pipelines copy-in / kernels / copy-out chunk by chunk: short chunks keep the
    # exposed first copy-in and last copy-out small, so it gets its own engine
    e2e_chunk = max(1, min(args.e2e_chunk, ne))
    heng = eng if e2e_chunk == chunk else Engine(library, H, W, max_batch=e2e_chunk, device=str(dev))
    heng.host_prepare(ne, want_all, not want_all)                       # allocations happen here, not in the call

    def host_call():
        if want_all:
            return heng.run_host(h_frames.numpy(), want_all=True, out_all=h_out.numpy())
        return heng.run_host(h_frames.numpy(), want_all=False, want_best=True, out_best=h_out.numpy())

    for _ in range(2):
        res = host_call()
    barrier()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        res = host_call()
    torch.cuda.synchronize(dev)
    e2e_s = time.perf_counter() - t0
    t = torch.tensor([e2e_s], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    e2e_value = world * ne * args.steps / float(t.item())
    h2d = ne * NPIX * 3
    d2h = ne * (S if want_all else 1) * NPIX * 3 + ne * (S * 9 * 8 + 4 + 4 + 12)
    # the host path must agree with the device path on the same frames: bytes and selection exactly; scores to 1e-9
    # (the two paths use different chunk sizes, hence different row chunks in the float64 guided-filter sums: the
    # float64 library's scores then differ in the last bits, the float32 library's not at all)
    d_scores = scores[:ne].cpu().numpy()
    score_diff = float(np.max(np.abs(res["scores"] - d_scores) / np.maximum(np.abs(d_scores), 1e-9)))
    same = score_diff <= 1e-9 and np.array_equal(res["best"], best[:ne].cpu().numpy())
    if want_all:
        same = same and np.array_equal(h_out.numpy()[0], out_all[0].cpu().numpy())
    else:
        same = same and np.array_equal(h_out.numpy()[:4], out_best[:4].cpu().numpy())
    if heng is not eng:
        heng.close()
    for e in engs:
        e.close()
    del d_frames, out_all, out_best, h_out, h_frames
    torch.cuda.empty_cache()

    peak, peak_src = peaks()
    alg = stage_alg_bytes(library)
    survey = survey_unit_bytes(library)
    traffic_all = load_traffic()
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
    calls_per_step = (B + chunk - 1) // chunk
    pipe_bytes = (3 + 3 * S) * NPIX if want_all else 6 * NPIX
    tr = traffic_all.get(library, {}) if isinstance(traffic_all.get(library), dict) else {}
    roofline = {
        "bound": "hbm", "kernel": dom["stage"], "achieved": dom["achieved_gbs"], "peak": peak, "unit": "GB/s",
        "frac": dom["achieved_gbs"] / peak,
        "traffic": tr.get(dom["stage"]), "traffic_source": traffic_all.get("_source"),
        "traffic_frames_per_call": traffic_all.get("_frames_per_call"),
        "peak_source": peak_src,
        "alg_bytes_per_launch": dom["alg_bytes_per_frame"] * chunk,
        "avg_launch_ms": dom["ms_per_step"] / calls_per_step,
        "share_of_step": dom["share"],
        "pipeline": {"alg_bytes_per_frame": pipe_bytes, "achieved": pipe_bytes * value / world / 1e9,
                     "frac": pipe_bytes * value / world / 1e9 / peak},
        "stages": table, "stage_timing": stage_note,
    }
    if dom["stage"] in survey:
        sb = survey[dom["stage"]] * NPIX
        roofline["survey_unit"] = {"what": "SURVEY.md 8(d) 'guided-filter dehaze (A given)' 6 N x 3 dehaze instances per frame",
                                   "alg_bytes_per_frame": sb,
                                   "achieved": dom["achieved_gbs"] * sb / dom["alg_bytes_per_frame"],
                                   "frac": dom["achieved_gbs"] * sb / dom["alg_bytes_per_frame"] / peak}
    return {
        "value": value, "ms_per_step": ms_max / args.steps, "launches": timed_launches, "clocks": clk.summary(),
        "roofline": roofline, "gather_checked": gather_ok, "S": S, "chunk": chunk, "lanes": lanes,
        "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                "frames_per_step": ne, "chunk": e2e_chunk, "api": "uwie_enhance_score_host (C ABI, pinned host buffers)",
                "matches_device_path": bool(same), "max_score_rel_diff_vs_device_path": score_diff},
    }


def content_check(torch, dist, dev, rank, world):
    """the same global batch sharded over `world` ranks gives, after the all-gather, the
    records one rank computes alone.  8 frames of 270x480 (seeds 900..907), image-index partition."""
    from underwater_image_enhancement_b200 import shard
    from underwater_image_enhancement_b200.engine import Engine
    from underwater_image_enhancement_b200.synth import synth_image
    M, h, w = 8, 270, 480
    lo, hi = shard.shard_range(M, world, rank)
    cap = shard.shard_capacity(M, world)
    eng = Engine("six", h, w, max_batch=M, device=str(dev))

    def run(seeds):
        if not len(seeds):
            z = torch.zeros
            return z((0, 6, 9), dtype=torch.float64, device=dev), z((0,), dtype=torch.int32, device=dev), z((0,), dtype=torch.int32, device=dev)
        fr = torch.from_numpy(np.stack([synth_image(900 + s, h, w) for s in seeds])).to(dev)
        r = eng.run(fr, want_all=False)
        eng.check()
        return r["scores"], r["best"], r["type"]

    sc, be, ty = run(list(range(lo, hi)))
    got = shard.unpack_records(shard.gather_records(shard.pack_records(sc, be, ty, cap)), 6)
    ok = True
    if rank == 0:
        full = run(list(range(M)))
        ok = all(torch.equal(a, b) for a, b in zip(got, full))
    eng.close()
    flag = torch.tensor([1 if ok else 0], device=dev)
    if world > 1:
        dist.all_reduce(flag, op=dist.ReduceOp.MIN)
    return bool(flag.item())


def run_native(args):
    import torch
    import torch.distributed as dist

    from underwater_image_enhancement_b200 import _lib
    from underwater_image_enhancement_b200.synth import synth_image

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise RuntimeError("bench.py (native arm) needs a CUDA device: there is no CPU fallback")
    numa = bind_to_gpu_numa(local, world) if (world > 1 and not args.no_numa_bind) else None
    print(f"[bench] rank {rank} numa binding: {numa}", file=sys.stderr, flush=True)
    torch.cuda.set_device(local)
    dev = torch.device(f"cuda:{local}")
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    _lib.load()                                                # fails loudly if the .so is missing

    distinct = min(args.distinct, args.batch)
    host = np.stack([synth_image(rank * distinct + i, H, W) for i in range(distinct)])
    main = measure_library(torch, dist, args, args.library, "all", dev, rank, world, host)
    p5 = None
    if not args.no_p5:
        p5 = measure_library(torch, dist, args, "es", "best", dev, rank, world, host)
    probe = pcie_probe(torch, dist, dev, world, args.probe_mb)
    same_on_n = content_check(torch, dist, dev, rank, world) if world > 1 else None

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return 0

    B, S = args.batch, main["S"]
    line = {
        "metric": METRIC, "value": main["value"], "unit": UNIT, "n_gpus": world, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": main["ms_per_step"], "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "u8+f32" if args.library == "six" else "u8+f64",
        "data": "synthetic",
        "config": {"workload": workload_name(args.library, B),
                   "library": args.library, "frames_per_gpu_per_step": B, "chunk": main["chunk"],
                   "distinct_frames": distinct, "strategies": S, "lanes": main["lanes"],
                   "l2": f"inputs {B * NPIX * 3 / 1e6:.0f} MB + outputs {B * S * NPIX * 3 / 1e6:.0f} MB per step exceed the 126 MB L2; no flush needed",
                   "parallelism": f"image-index sharding x{world}, NCCL all-gather of score records",
                   "numa_bind": numa},
        "e2e": main["e2e"],
        "gpu_launches": main["launches"],
        "clocks": main["clocks"],
        "roofline": main["roofline"],
    }
    if probe:
        line["e2e"]["pcie_probe"] = probe
    if world > 1:
        line["gather_checked"] = bool(main["gather_checked"] and same_on_n)
        line["gather_check"] = {"bench_records_in_rank_order": main["gather_checked"],
                                "same_batch_sharded_equals_one_rank": same_on_n}
    if p5:
        line["p5"] = {"workload": workload_name("es", B) + ", best-only contract of main.py:122-155 (6 N bytes per frame)",
                      "value": p5["value"], "unit": UNIT, "ms_per_step": p5["ms_per_step"], "dtype": "u8+f64",
                      "e2e": p5["e2e"], "gpu_launches": p5["launches"], "roofline": p5["roofline"],
                      "gather_checked": p5["gather_checked"]}
    if world == 1 and not args.no_cpu_baseline:
        kind = cpu_kind()
        workers = host_cores()
        per = max(1, args.cpu_frames // workers) if args.cpu_frames else 1
        v, busy = cpu_throughput(args.library, workers, per, kind=kind)
        line["cpu_baseline"] = {"value": v, "unit": UNIT, "cores": workers, "kind": kind,
                                "sample": cpu_sample_text(kind, workers, per) + f" ({busy:.1f} s)"}
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
               "vgg_preprocess": o.vgg_preprocess, "features79": o.fe_all, "quality_scores": o.qa_total,
               "oracle": o}
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
