"""Seeded synthetic inputs of the shapes BASELINE.json names. Host-side numpy only — inputs,
not compute. Generators follow the reference's order "base first, then queries" (src/dataset_loader.h:19-24).

  gauss : iid N(0,1) per component, un-normalised — the reference's own generator (src/randomgeometry.h:87-95)
  sift  : u8-valued floats. variant "iid" uniform_int[0,255]; variant "lowrank" = clamp(round(128+48*(Az)/sqrt(r)))
  deep  : low-rank Gaussian (r=24), L2-normalised, d=96
  gist  : low-rank (r=32) mapped to [0,1], d=960
"""
from __future__ import annotations

import numpy as np


def _lowrank(rng, n, d, r, chunk=1 << 18):
    A = np.random.default_rng(1).standard_normal((d, r)).astype(np.float32)
    out = np.empty((n, d), np.float32)
    for i in range(0, n, chunk):
        z = rng.standard_normal((min(chunk, n - i), r), dtype=np.float32)
        out[i:i + chunk] = z @ A.T
    return out / np.float32(np.sqrt(r))


def make(kind: str, n: int, m: int, d: int | None = None, seed: int | None = None, variant: str = "lowrank"):
    """returns (base [n][d] float32, queries [m][d] float32)"""
    if kind == "gauss":
        d = d or 128
        rng = np.random.default_rng(12345 if seed is None else seed)
        x = rng.standard_normal((n + m, d), dtype=np.float32)
    elif kind == "sift":
        d = d or 128
        rng = np.random.default_rng(2024 if seed is None else seed)
        if variant == "iid":
            x = rng.integers(0, 256, size=(n + m, d), dtype=np.uint8).astype(np.float32)
        else:
            x = _lowrank(rng, n + m, d, 16)
            x = np.clip(np.rint(128.0 + 48.0 * x), 0, 255).astype(np.float32)
    elif kind == "deep":
        d = d or 96
        rng = np.random.default_rng(96 if seed is None else seed)
        x = _lowrank(rng, n + m, d, 24)
        x /= np.maximum(np.linalg.norm(x, axis=1, keepdims=True), 1e-12).astype(np.float32)
    elif kind == "gist":
        d = d or 960
        rng = np.random.default_rng(960 if seed is None else seed)
        x = _lowrank(rng, n + m, d, 32)
        x = np.clip(0.5 + 0.15 * x, 0.0, 1.0).astype(np.float32)
    else:
        raise ValueError(kind)
    x = np.ascontiguousarray(x, dtype=np.float32)
    return x[:n], x[n:]
