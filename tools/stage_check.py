"""Stage-by-stage comparison of the CUDA path with the oracle on small seeded frames (GPU box).

    python tools/stage_check.py [--lib six|es] [--cases 0:120:160,7:250:333]

Prints one line per stage and frame; used while bringing kernels up.  The pytest parity suite
(tests/test_gpu_parity.py) is the gate; this is the microscope.
"""
from __future__ import annotations

import argparse
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from oracle import restatement as o                                    # noqa: E402
from underwater_image_enhancement_b200.engine import Engine            # noqa: E402
from underwater_image_enhancement_b200.synth import synth_image        # noqa: E402


def report(tag, ok, extra=""):
    print(f"  [{'ok ' if ok else 'BAD'}] {tag} {extra}", flush=True)
    return ok


def check_six(seed, h, w):
    print(f"== six seed={seed} {h}x{w}", flush=True)
    frame = synth_image(seed, h, w)
    eng = Engine("six", h, w, max_batch=2)
    fr = torch.from_numpy(np.stack([frame, frame])).cuda()
    res = eng.run(fr)
    eng.check()
    img = o.to_unit_f32(frame)
    kind = o.ss_image_type(img)
    cur = o.ss_correct_cast(img, kind)
    good = True
    good &= report("type", int(res["type"][0]) == o.IMAGE_TYPES.index(kind), f"{int(res['type'][0])} vs {kind}")
    trace = []
    light = o.quadtree_airlight(cur, 1, trace)
    qt = eng.debug("q_trace", np.float64, (2, 24, 4))[0]
    for lv, (_, _, _, _, q) in enumerate(trace):
        d = np.abs(qt[lv] - np.array(q)).max()
        same = int(np.argmax(qt[lv])) == int(np.argmax(q))
        if not same or d > 1e-5:
            good &= report(f"quad level {lv}", same, f"maxdiff {d:.3e} gpu {qt[lv]} ref {q}")
    good &= report("light", np.array_equal(res["light"][0].cpu().numpy(), light), f"{res['light'][0].cpu().numpy()} vs {light}")
    # recovered frames of the three dehaze strategies
    rec = eng.debug("recovered", np.float32, (2, 3, h, w, 3))[0]
    for k, (omega, r, eps) in enumerate(((0.3, 20, 0.5), (0.5, 15, 0.5), (0.7, 10, 0.1))):
        t = o.ss_transmission(cur, light, omega, r, eps)
        ref = o.ss_restore(cur, light, t)
        d = np.abs(rec[k].astype(np.float64) - ref.astype(np.float64))
        good &= report(f"recovered s{k + 1}", d.max() < 1e-6, f"maxdiff {d.max():.3e} n_neq {(rec[k] != ref).sum()}")
    out = res["out"][0].cpu().numpy()
    ref = o.p6_frame(frame)
    for s in range(6):
        neq = (out[s] != ref["out"][s])
        dmax = np.abs(out[s].astype(int) - ref["out"][s].astype(int)).max()
        good &= report(f"out s{s + 1}", dmax <= 1, f"bytes differing {neq.sum()} / {neq.size}, max |d| {dmax}")
    sc = res["scores"][0].cpu().numpy()
    for s in range(6):
        rel = np.abs(sc[s] - ref["scores"][s]) / np.maximum(np.abs(ref["scores"][s]), 1e-9)
        good &= report(f"scores s{s + 1}", rel.max() < 1e-4, f"max rel {rel.max():.2e} gpu {np.round(sc[s], 4)} ref {np.round(ref['scores'][s], 4)}")
    good &= report("best", int(res["best"][0]) == ref["best"], f"{int(res['best'][0])} vs {ref['best']}")
    same01 = all(torch.equal(res[k][0], res[k][1]) for k in ("out", "scores", "best", "type", "light"))
    good &= report("batch slot 0 == slot 1", same01)
    eng.close()
    return good


def check_es(seed, h, w):
    print(f"== es seed={seed} {h}x{w}", flush=True)
    frame = synth_image(seed, h, w)
    eng = Engine("es", h, w, max_batch=2)
    fr = torch.from_numpy(np.stack([frame, frame])).cuda()
    res = eng.run(fr, want_best=True)
    eng.check()
    ref = o.p5_frame(frame)
    good = True
    good &= report("light", np.array_equal(res["light"][0].cpu().numpy(), ref["light"]), f"{res['light'][0].cpu().numpy()} vs {ref['light']}")
    out = res["out"][0].cpu().numpy()
    for s in range(5):
        neq = (out[s] != ref["out"][s])
        dmax = np.abs(out[s].astype(int) - ref["out"][s].astype(int)).max()
        good &= report(f"out {s}", dmax <= 1, f"bytes differing {neq.sum()} / {neq.size}, max |d| {dmax}")
    sc = res["scores"][0].cpu().numpy()
    for s in range(5):
        rel = np.abs(sc[s] - ref["scores"][s]) / np.maximum(np.abs(ref["scores"][s]), 1e-9)
        good &= report(f"scores {s}", rel.max() < 1e-4, f"max rel {rel.max():.2e} gpu {np.round(sc[s], 4)} ref {np.round(ref['scores'][s], 4)}")
    good &= report("best", int(res["best"][0]) == ref["best"], f"{int(res['best'][0])} vs {ref['best']}")
    good &= report("best image", np.array_equal(res["best_out"][0].cpu().numpy(), out[int(res["best"][0])]))
    eng.close()
    return good


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--lib", default="six,es")
    ap.add_argument("--cases", default="0:120:160,1:120:160,2:120:160,3:135:241,7:250:333")
    a = ap.parse_args()
    ok = True
    for case in a.cases.split(","):
        seed, h, w = (int(x) for x in case.split(":"))
        if "six" in a.lib:
            ok &= check_six(seed, h, w)
        if "es" in a.lib:
            ok &= check_es(seed, h, w)
    print("ALL OK" if ok else "SOME STAGES DIFFER")
    return 0 if ok else 1


if __name__ == "__main__":
    sys.exit(main())
