"""Synthetic inputs of SURVEY.md §8(d): random triangle soup M(N) and incoherent rays.

numpy twin of oracle/vo_synth.h (counter-based splitmix64; fp32 in [0,1) = (x >> 40) * 2^-24).
Used by bench.py and the tests; generating inputs is not part of the timed path.
"""
from __future__ import annotations

import numpy as np

_K = np.uint64(0xD1342543DE82EF95)
_G = np.uint64(0x9E3779B97F4A7C15)
_M1 = np.uint64(0xBF58476D1CE4E5B9)
_M2 = np.uint64(0x94D049BB133111EB)


def splitmix64(seed: int, k: np.ndarray) -> np.ndarray:
    with np.errstate(over="ignore"):
        z = np.uint64(seed) * _K + (k.astype(np.uint64) + np.uint64(1)) * _G
        z = (z ^ (z >> np.uint64(30))) * _M1
        z = (z ^ (z >> np.uint64(27))) * _M2
        return z ^ (z >> np.uint64(31))


def rnd01(seed: int, k: np.ndarray) -> np.ndarray:
    return (splitmix64(seed, k) >> np.uint64(40)).astype(np.float32) * np.float32(2.0 ** -24)


def _tri_block(seed: int, n_total: int, idx: np.ndarray) -> np.ndarray:
    r = np.float32(2.0 * (1.0 / float(n_total)) ** (1.0 / 3.0))
    k = idx.astype(np.uint64)[:, None] * np.uint64(12) + np.arange(12, dtype=np.uint64)[None, :]
    u = rnd01(seed, k)                                   # (m, 12)
    c = np.float32(2.0) * u[:, 0:3] - np.float32(1.0)
    xi = u[:, 3:12].reshape(-1, 3, 3)
    off = r * (xi - np.float32(0.5))
    return (c[:, None, :] + off).astype(np.float32).reshape(-1, 9)


def synth_mesh(n: int, seed: int = 1, chunk: int = 1 << 20) -> np.ndarray:
    """(n, 9) float32 triangle soup; degenerate triangles (area < 1e-12) are redrawn with index + N."""
    out = np.empty((n, 9), np.float32)
    for b in range(0, n, chunk):
        e = min(n, b + chunk)
        idx = np.arange(b, e, dtype=np.uint64)
        blk = _tri_block(seed, n, idx)
        while True:
            v0, v1, v2 = blk[:, 0:3], blk[:, 3:6], blk[:, 6:9]
            e1, e2 = v1 - v0, v2 - v0
            cx = e1[:, 1] * e2[:, 2] - e1[:, 2] * e2[:, 1]
            cy = e1[:, 2] * e2[:, 0] - e1[:, 0] * e2[:, 2]
            cz = e1[:, 0] * e2[:, 1] - e1[:, 1] * e2[:, 0]
            a2 = cx * cx + cy * cy + cz * cz
            bad = np.nonzero(~(a2 >= np.float32(4e-24)))[0]
            if bad.size == 0:
                break
            idx[bad] += np.uint64(n)
            blk[bad] = _tri_block(seed, n, idx[bad])
        out[b:e] = blk
    return out


def synth_rays(n: int, seed: int = 2, seed_any: int = 3, chunk: int = 1 << 21) -> dict:
    """SoA incoherent rays: origins uniform in [-1,1]^3, directions uniform on the sphere
    (square_to_sphere, reference src/math/warp.h:72-78). 'tmax' = 1e30 (closest set);
    'tmax_any' uniform in [0.05, 0.5] (any-hit set)."""
    r = {f: np.empty(n, np.float32) for f in ("ox", "oy", "oz", "tmin", "dx", "dy", "dz", "tmax", "tmax_any")}
    for b in range(0, n, chunk):
        e = min(n, b + chunk)
        i = np.arange(b, e, dtype=np.uint64)
        k = i * np.uint64(8)
        for a, f in enumerate(("ox", "oy", "oz")):
            r[f][b:e] = np.float32(2.0) * rnd01(seed, k + np.uint64(a)) - np.float32(1.0)
        u0, u1 = rnd01(seed, k + np.uint64(3)), rnd01(seed, k + np.uint64(4))
        z = np.float32(1.0) - np.float32(2.0) * u0
        rr = np.sqrt(np.maximum(np.float32(0.0), np.float32(1.0) - z * z)).astype(np.float32)
        phi = 2.0 * np.pi * u1.astype(np.float64)
        r["dx"][b:e] = (rr.astype(np.float64) * np.cos(phi)).astype(np.float32)
        r["dy"][b:e] = (rr.astype(np.float64) * np.sin(phi)).astype(np.float32)
        r["dz"][b:e] = z
        r["tmin"][b:e] = 0.0
        r["tmax"][b:e] = 1e30
        r["tmax_any"][b:e] = np.float32(0.05) + np.float32(0.45) * rnd01(seed_any, i)
    return r
