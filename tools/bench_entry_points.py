"""Device timing of the entry points beside the fused six-strategy pipeline (GPU box), one JSON line each:

    python tools/bench_entry_points.py [--out gpurun_out/entry_points.json]
    python bench.py --entry-points [--out ...]        # the same, with the CPU restatement timed beside each line

  config2_clahe / config2_dehaze   BASELINE.json configs[2]: CLAHE only / dark-channel dehaze only on 3840x2160
                                   frames (engine with ONE strategy of enhancement_strategies.py)
  vgg_enhance / vgg_features / vgg_preprocess
                                   BASELINE.json configs[3] pieces (use_trained_model.py:55, 39-46, 103) at
                                   batch 64 of 1080p
  features79                       feature_extraction.py:249-297 at 1080p
  quality_scores                   quality_assessment.py:215-286 at 1080p

Frames are resident in HBM (this is the device number, like bench.py's `value`); every input batch is larger
than the 126 MB L2.  `achieved` = SURVEY.md 8(d) algorithmic bytes per frame x frames / CUDA-event time;
`peak` = MEASURED_PEAKS.json hbm_gbs.  This file never touches oracle/: bench.py's cpu_baseline leg passes in
the CPU callables (`cpu`), each timed on ONE frame of the same workload on one host core.
"""
from __future__ import annotations

import argparse
import json
import os
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from underwater_image_enhancement_b200 import compat, features, prims, quality, vgg      # noqa: E402
from underwater_image_enhancement_b200.engine import Engine                       # noqa: E402
from underwater_image_enhancement_b200.synth import synth_image                   # noqa: E402


def peak_gbs():
    try:
        return float(json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"]), "measured"
    except Exception:                                                              # noqa: BLE001
        return 6500.0, "fallback (B200_PROFILING.md)"


def time_ms(fn, warmup=3, reps=5):
    for _ in range(warmup):
        fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(reps):
        fn()
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) / reps


def tiled(h, w, distinct, batch, seed0):
    base = np.stack([synth_image(seed0 + i, h, w) for i in range(distinct)])
    return np.concatenate([base] * ((batch + distinct - 1) // distinct))[:batch]


def cpu_time(cpu, name, *a):
    if not cpu or name not in cpu:
        return None
    t0 = time.perf_counter()
    cpu[name](*a)
    return time.perf_counter() - t0


def main(cpu=None, out_path=None):
    if out_path is None:
        ap = argparse.ArgumentParser()
        ap.add_argument("--out", default="")
        out_path = ap.parse_args().out
    if not torch.cuda.is_available():
        raise SystemExit("needs a CUDA device: the entry points have no CPU path")
    peak, peak_src = peak_gbs()
    lines = []

    def emit(name, ref, h, w, batch, bytes_per_px, ms, dtype, cpu_s=None, note="", alg_bytes=None):
        n = h * w
        alg = bytes_per_px * n if alg_bytes is None else alg_bytes
        gbs = alg * batch / (ms * 1e-3) / 1e9
        rec = {"entry": name, "reference": ref, "size": f"{w}x{h}", "batch": batch, "dtype": dtype,
               "ms_per_batch": round(ms, 4), "frames_per_s": round(batch / (ms * 1e-3), 1),
               "alg_bytes_per_frame": alg,
               "roofline": {"bound": "hbm", "achieved": round(gbs, 1), "peak": peak, "unit": "GB/s",
                            "frac": round(gbs / peak, 4), "peak_source": peak_src}}
        if cpu_s is not None:
            rec["cpu_baseline"] = {"value": round(1.0 / cpu_s, 4), "unit": "frames/s", "cores": 1, "kind": "port",
                                   "sample": "1 frame of the same workload (bench.py cpu_baseline leg)"}
        if note:
            rec["note"] = note
        lines.append(rec)
        print(json.dumps(rec), flush=True)

    # ---- configs[2]: the ISOLATED bandwidth-bound kernels on 3840x2160 frames (2 N and 7 N)
    h, w, batch = 2160, 3840, 32                      # 32 planes in + out = 531 MB, 32 frames + t = 1.86 GB: beyond the L2
    import cv2
    base = tiled(h, w, 4, 4, 50)
    lplanes = np.stack([cv2.cvtColor(f, cv2.COLOR_RGB2LAB)[:, :, 0] for f in base])
    planes = torch.from_numpy(np.concatenate([lplanes] * (batch // 4))).cuda()
    pout = torch.empty_like(planes)
    ws = torch.empty(prims._lib.load().uwie_prim_clahe_lab_workspace_bytes(batch, 8, 8) + 256, dtype=torch.uint8, device="cuda")
    ms = time_ms(lambda: prims.clahe_lab(planes, 2.0, (8, 8), out=pout, workspace=ws))
    cpu_s = None
    if cpu and "oracle" in cpu:
        op = cv2.createCLAHE(clipLimit=2.0, tileGridSize=(8, 8))
        cv2.setNumThreads(1)
        t0 = time.perf_counter(); op.apply(lplanes[0]); cpu_s = time.perf_counter() - t0
    emit("prim_clahe_lab", "enhancement_strategies.py:302-303 / six_stadigy.py:205-206 (cv2 CLAHE on the Lab L plane)",
         h, w, batch, 2, ms, "u8", cpu_s, "isolated kernel pair (tile histograms + LUTs, interpolation): u8 L plane in, u8 L plane out; "
         "bit-exact vs cv2; CPU side = cv2.createCLAHE().apply on one plane, one thread")
    del planes, pout
    frames = torch.from_numpy(np.concatenate([base] * (batch // 4))).cuda()
    light = torch.tensor([[0.31, 0.62, 0.71]] * batch, dtype=torch.float32, device="cuda")
    tout = torch.empty((batch, h, w), dtype=torch.float32, device="cuda")
    ms = time_ms(lambda: prims.transmission_head_u8(frames, light, 0.5, 1e-6, True, out=tout))
    cpu_s = None
    if cpu and "oracle" in cpu:
        img0 = base[0].astype(np.float32) / 255.0
        t0 = time.perf_counter()
        np.clip(1 - np.float32(0.5) * np.min(img0 / (light[0].cpu().numpy().reshape(1, 1, 3) + 1e-6), axis=2), 0.1, 1.0)
        cpu_s = time.perf_counter() - t0
    emit("prim_transmission_head", "six_stadigy.py:170-174 / enhancement_strategies.py:221-225 (dark channel -> raw transmission)",
         h, w, batch, 7, ms, "u8->f32", cpu_s, "isolated kernel, TMA-staged tiles (cp.async.bulk + mbarrier): u8 RGB in, float32 t plane out; "
         "bit-exact vs NumPy float32")
    del tout, frames

    # ---- configs[2] as whole strategies (engine with ONE strategy + scoring of its output)
    batch = 8
    frames = torch.from_numpy(tiled(h, w, 4, batch, 50)).cuda()
    for name, sname, params, ref in (
            ("config2_clahe", "clahe_enhancement", {"clip_limit": 2.0, "tile_grid_size": (8, 8)},
             "enhancement_strategies.py:400-420"),
            ("config2_dehaze", "light_enhancement", {"omega": 0.4, "guided_radius": 10, "L_low": 15, "L_high": 95},
             "enhancement_strategies.py:423-458")):
        eng = Engine("es", h, w, max_batch=batch, strategies=[compat._es_strategy(sname, params)])
        out = torch.empty((batch, 1, h, w, 3), dtype=torch.uint8, device="cuda")
        ms = time_ms(lambda: eng.run(frames, want_all=True, out_all=out))
        eng.check()
        eng.close()
        cpu_s = cpu_time(cpu, "es_strategy_scored", frames[0].cpu().numpy().astype(np.float32) / 255.0, sname, params)
        emit(name, ref, h, w, batch, 6, ms, "u8+f64", cpu_s, "strategy + scoring of its output; u8 RGB in, u8 RGB out")
    del frames

    # ---- configs[3] pieces at batch 64 of 1080p
    h, w, batch = 1080, 1920, 64
    fr_np = tiled(h, w, 8, batch, 60)
    frames = torch.from_numpy(fr_np).cuda()
    img = (frames.to(torch.float32) / 255.0).permute(0, 3, 1, 2).contiguous()
    rng = np.random.default_rng(0)
    params = {"L_low": rng.uniform(1, 15, batch).tolist(), "L_high": rng.uniform(85, 99, batch).tolist(),
              "omega": rng.uniform(0.3, 0.9, batch).tolist(), "gamma": rng.uniform(0.8, 1.4, batch).tolist()}
    enh = vgg.DifferentiableEnhancement()
    ms = time_ms(lambda: enh(img, params))
    cpu_s = cpu_time(cpu, "vgg_enhance", img[:1].cpu().numpy(), params["L_low"][:1], params["L_high"][:1],
                     params["omega"][:1], params["gamma"][:1])
    emit("vgg_enhance", "vgg_16_UIE.py:32-128", h, w, batch, 24, ms, "f32", cpu_s, "f32 NCHW in and out")
    pdev = torch.tensor([params["L_low"], params["L_high"], params["omega"], params["gamma"]], dtype=torch.float32, device="cuda")
    ms = time_ms(lambda: vgg.enhance_frames_u8(frames, pdev, True))
    emit("vgg_enhance_u8", "vgg_16_UIE.py:32-128 on imread bytes (use_trained_model.py:119)", h, w, batch, 15, ms, "u8->f32", None,
         "u8 HWC in, f32 NCHW out (15 N): order statistics from the byte histograms, frame read twice as bytes")
    ms = time_ms(lambda: vgg.extract_all_features(frames))
    cpu_s = cpu_time(cpu, "vgg_features", fr_np[0])
    emit("vgg_features", "vgg_16_UIE.py:435-466", h, w, batch, 3, ms, "u8->f32", cpu_s)
    ms = time_ms(lambda: vgg.preprocess_for_vgg(frames))
    cpu_s = cpu_time(cpu, "vgg_preprocess", fr_np[0])
    emit("vgg_preprocess", "use_trained_model.py:39-46", h, w, batch, 3, ms, "u8->f32", cpu_s,
         "reads 4 byte taps per value of the 224x224x3 network input, not the whole frame: algorithmic bytes = "
         "224*224*3*(4 taps + 4 bytes of float32 output)", alg_bytes=224 * 224 * 3 * 8)
    del img

    # ---- configs[3] END TO END at batch 64 of 1080p: features + 224^2 input -> parameter network (the reference's own
    # ImprovedVGGParameterNet, random-init, seed 0, stock PyTorch) -> clamps + enhancement with device-resident parameters
    try:
        from underwater_image_enhancement_b200.predictor import EnhancementPredictor, load_network_class
        torch.manual_seed(0)
        net = load_network_class()(pretrained=False, hidden_dim=256, use_features=True)
        for dt, tag in ((torch.float32, "fp32 network"), (torch.bfloat16, "bf16 autocast network")):
            pred = EnhancementPredictor(model=net, net_dtype=dt)
            ms = time_ms(lambda: pred.enhance_batch(frames), warmup=2, reps=3)
            ms_net = time_ms(lambda: pred.predict_parameters_batch(frames), warmup=2, reps=3)
            emit("config3_vgg_guided_e2e", "use_trained_model.py:53-111 (predict_parameters + enhance_image), vgg_16_UIE.py:209-250",
                 h, w, batch, 3 + 12, ms, "u8->f32", None,
                 f"{tag}; uint8 frames resident in HBM -> float32 NCHW enhanced (15 N bytes per frame: the uint8->float32 NCHW "
                 f"conversion of use_trained_model.py:48-51 is a torch op inside the timed region); predict_parameters alone "
                 f"{ms_net:.2f} ms per batch; no per-image host round trip")
    except Exception as e:                                                          # noqa: BLE001
        print(json.dumps({"entry": "config3_vgg_guided_e2e", "unavailable": f"{type(e).__name__}: {e}"}), flush=True)

    # ---- FE 79 features and stand-alone QA at 1080p
    batch = 32
    sub = frames[:batch].contiguous()
    ms = time_ms(lambda: features.extract_all_features_batch(sub), warmup=2, reps=3)
    cpu_s = cpu_time(cpu, "features79", fr_np[0].astype(np.float32) / 255.0)
    emit("features79", "feature_extraction.py:249-297", h, w, batch, 3, ms, "u8+f32/f64", cpu_s,
         "includes the full-frame DCT (two float32 GEMMs, 12.4 GFLOP per frame): compute-bound piece, SURVEY 8(d)")
    ms = time_ms(lambda: quality.quality_scores_batch(sub))
    cpu_s = cpu_time(cpu, "quality_scores", fr_np[0].astype(np.float32) / 255.0)
    emit("quality_scores", "quality_assessment.py:215-286", h, w, batch, 3, ms, "u8", cpu_s)

    # ---- (f-1) the file loop: PNG files in -> six PNG files out per frame, through the pinned double-buffered reader / writer
    try:
        import shutil
        import tempfile
        import cv2
        from underwater_image_enhancement_b200 import compat as _compat
        tmp = tempfile.mkdtemp(prefix="uwie_io_", dir="/dev/shm" if os.path.isdir("/dev/shm") else None)
        src, dst = os.path.join(tmp, "in"), os.path.join(tmp, "out")
        os.makedirs(src)
        nfiles = 48
        for i in range(nfiles):
            cv2.imwrite(os.path.join(src, f"f{i:03d}.png"), cv2.cvtColor(fr_np[i % len(fr_np)], cv2.COLOR_RGB2BGR))
        workers = max(2, min(32, (os.cpu_count() or 4)))
        _compat.process_all_images_all_strategies(src, dst, batch=8, workers=workers, write_csv=False)      # warm-up (engine, pools)
        shutil.rmtree(dst)
        t0 = time.perf_counter()
        log = _compat.process_all_images_all_strategies(src, dst, batch=8, workers=workers)
        dt = time.perf_counter() - t0
        t1 = time.perf_counter()
        for i in range(8):
            cv2.imread(os.path.join(src, f"f{i:03d}.png"))
        dec = (time.perf_counter() - t1) / 8
        t1 = time.perf_counter()
        for i in range(8):
            cv2.imwrite(os.path.join(tmp, f"e{i}.png"), fr_np[i % len(fr_np)])
        enc = (time.perf_counter() - t1) / 8
        ok = sum(r["status"] == "success" for r in log)
        rec = {"entry": "file_loop_six", "reference": "six_stadigy.py:326-470 (process_all_images_all_strategies)", "size": f"{w}x{h}",
               "files": nfiles, "outputs_written": ok, "seconds": round(dt, 3), "frames_per_s": round(nfiles / dt, 1),
               "workers": workers, "png_decode_ms_one_thread": round(dec * 1e3, 1), "png_encode_ms_one_thread": round(enc * 1e3, 1),
               "note": "PNG files on tmpfs in, 6 PNG files + processing_log.csv out; decode / encode on worker threads (cv2 releases the "
                       "GIL), pinned double-buffered slots around uwie_enhance_score_host; the ceiling is the host's PNG codec: "
                       "workers / (decode + 6 x encode) frames/s"}
        lines.append(rec)
        print(json.dumps(rec), flush=True)
        _compat.close_engines()
        shutil.rmtree(tmp, ignore_errors=True)
    except Exception as e:                                                          # noqa: BLE001
        print(json.dumps({"entry": "file_loop_six", "unavailable": f"{type(e).__name__}: {e}"}), flush=True)

    if out_path:
        with open(out_path, "w") as f:
            for rec in lines:
                f.write(json.dumps(rec) + "\n")


if __name__ == "__main__":
    main()
