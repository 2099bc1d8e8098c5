"""Decode / encode front end of the file loops: main.py:88-108, 150-155,
six_stadigy.py:392-406, 430-447, use_trained_model.py:115-140.

At device speed the reference's loop body (`cv2.imread` -> work -> `cv2.imwrite`, one file at a time) is all
file I/O.  Here decode, device work and encode overlap:

* `FrameReader`  worker threads decode files (cv2 releases the GIL) ahead of the consumer; frames of one size
  are packed into one of two PINNED host slots, so the engine's host entry point copies them to the device
  asynchronously while the next slot fills;
* `FrameWriter`  the enhanced frames land in one of two pinned output slots; worker threads encode and write
  them while the next batch runs.  A slot is reused only when all its writes have finished.

Nothing here computes on pixels: BGR<->RGB is a channel order fix of the codec's layout, done by OpenCV on the
host next to the codec, exactly where the reference does it (main.py:105, six_stadigy.py:405-406, 433).
"""
from __future__ import annotations

from collections import deque
from concurrent.futures import ThreadPoolExecutor
from pathlib import Path

import numpy as np
import torch

IMAGE_SUFFIXES = (".jpg", ".jpeg", ".png", ".tif", ".tiff")          # six_stadigy.py:361


def list_images(folder, suffixes=IMAGE_SUFFIXES):
    """Files of `folder` with one of `suffixes` (case-insensitive), sorted by name."""
    return sorted(p for p in Path(folder).iterdir() if p.is_file() and p.suffix.lower() in suffixes)


def decode_rgb(path):
    """cv2.imread(path) (default flags: always 8-bit, 3 channels, as in main.py:91 and six_stadigy.py:398) and the
    BGR -> RGB channel fix of main.py:105 / six_stadigy.py:406: uint8 RGB (H, W, 3), or None when unreadable."""
    import cv2
    img = cv2.imread(str(path))
    if img is None:
        return None
    return cv2.cvtColor(img, cv2.COLOR_BGR2RGB)


def _pinned(shape):
    return torch.empty(shape, dtype=torch.uint8, pin_memory=torch.cuda.is_available())


class FrameReader:
    """Iterate over batches `(paths, frames)`; `frames` is a uint8 NumPy view [n, H, W, 3] of a pinned slot that stays
    valid until the batch after the next is requested.  Files that cannot be decoded are reported through
    `failed` (list of paths) and skipped, like the reference's `continue` (main.py:92-94, six_stadigy.py:399-403).
    A batch never mixes sizes: it ends early when the next file has another shape."""

    def __init__(self, files, batch=8, workers=4, prefetch=None):
        self.files = [Path(f) for f in files]
        self.batch = max(1, int(batch))
        self.failed = []
        self._pool = ThreadPoolExecutor(max_workers=max(1, workers), thread_name_prefix="uwie-decode")
        self._prefetch = prefetch or 2 * self.batch
        self._slots = {}                                     # (H, W) -> [pinned tensors], used in turn
        self._turn = {}

    def _slot(self, h, w):
        key = (h, w)
        if key not in self._slots:
            self._slots[key] = [_pinned((self.batch, h, w, 3)) for _ in range(2)]
            self._turn[key] = 0
        t = self._slots[key][self._turn[key]]
        self._turn[key] ^= 1
        return t

    def __iter__(self):
        pending = deque()
        it = iter(self.files)

        def top_up():
            while len(pending) < self._prefetch:
                p = next(it, None)
                if p is None:
                    return
                pending.append((p, self._pool.submit(decode_rgb, p)))

        top_up()
        carry = None                                         # a decoded frame that did not fit the previous batch
        while pending or carry is not None:
            paths, frames, shape = [], [], None
            while len(paths) < self.batch:
                if carry is not None:
                    p, img = carry
                    carry = None
                elif pending:
                    p, fut = pending.popleft()
                    top_up()
                    try:
                        img = fut.result()
                    except Exception:                        # noqa: BLE001
                        img = None
                    if img is None:
                        self.failed.append(p)
                        continue
                else:
                    break
                if shape is None:
                    shape = img.shape[:2]
                elif img.shape[:2] != shape:
                    carry = (p, img)
                    break
                paths.append(p)
                frames.append(img)
            if not paths:
                continue
            slot = self._slot(*shape).numpy()
            for i, f in enumerate(frames):
                np.copyto(slot[i], f)
            yield paths, slot[:len(paths)]

    def close(self):
        self._pool.shutdown(wait=False, cancel_futures=True)


class FrameWriter:
    """Pinned output slots + encode threads.  `slot(shape)` hands out a NumPy view of a pinned buffer whose previous
    writes have all finished; `write(path, rgb)` queues one `cv2.imwrite` (RGB -> BGR like six_stadigy.py:433) and
    returns a future with cv2's boolean."""

    def __init__(self, workers=4):
        self._pool = ThreadPoolExecutor(max_workers=max(1, workers), thread_name_prefix="uwie-encode")
        self._slots = {}
        self._turn = {}
        self._busy = {}

    def slot(self, shape):
        key = tuple(int(v) for v in shape)
        if key not in self._slots:
            self._slots[key] = [_pinned(key) for _ in range(2)]
            self._turn[key] = 0
            self._busy[key] = [[], []]
        k = self._turn[key]
        self._turn[key] ^= 1
        for fut in self._busy[key][k]:
            fut.result()
        self._busy[key][k] = []
        self._current = (key, k)
        return self._slots[key][k].numpy()

    @staticmethod
    def _encode(path, rgb):
        import cv2
        return bool(cv2.imwrite(str(path), cv2.cvtColor(rgb, cv2.COLOR_RGB2BGR)))

    def write(self, path, rgb):
        fut = self._pool.submit(self._encode, path, rgb)
        key, k = self._current
        self._busy[key][k].append(fut)
        return fut

    def drain(self):
        for key in self._busy:
            for k in (0, 1):
                for fut in self._busy[key][k]:
                    fut.result()
                self._busy[key][k] = []

    def close(self):
        self.drain()
        self._pool.shutdown(wait=True)
