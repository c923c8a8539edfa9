"""Helpers for the reference-held parity pins (tests/golden/reference_renders/, tools/import_reference_renders.py).

Two measurements against an image the REFERENCE rendered:
  * silhouettes: our primary-hit id map (instance, face) gives the pixel positions of every object boundary; the reference
    image's gradient magnitude, averaged over those positions shifted by (dy, dx), must peak at (0, 0) - in the whole frame and in
    each quadrant (a wrong field of view, aspect, handedness, matrix convention or vertex order moves the quadrants differently);
  * radiometry: block means of the display-referred image (exposure -> tone map -> sRGB OETF, frame_buffer.cpp:72-74,135-154)
    and per-wall mean colours, in 8-bit sRGB units."""
from pathlib import Path

import numpy as np

GOLDEN = Path(__file__).resolve().parent / "golden" / "reference_renders"


def load_png(path) -> np.ndarray:
    from PIL import Image
    return np.asarray(Image.open(path).convert("RGB")).astype(np.float64)


def _blur3(a):
    """separable [1 2 1]/4 twice (about a sigma-1 gaussian), edges clamped; a is (h, w, c)"""
    for _ in range(2):
        p = np.pad(a, ((1, 1), (0, 0), (0, 0)), mode="edge")
        a = 0.25 * p[:-2] + 0.5 * p[1:-1] + 0.25 * p[2:]
        p = np.pad(a, ((0, 0), (1, 1), (0, 0)), mode="edge")
        a = 0.25 * p[:, :-2] + 0.5 * p[:, 1:-1] + 0.25 * p[:, 2:]
    return a


def grad_mag(img) -> np.ndarray:
    sm = _blur3(img.astype(np.float64))
    gx = np.zeros(sm.shape[:2]); gy = np.zeros(sm.shape[:2])
    gx[:, 1:-1] = np.abs(sm[:, 2:] - sm[:, :-2]).sum(-1)
    gy[1:-1, :] = np.abs(sm[2:, :] - sm[:-2, :]).sum(-1)
    return np.hypot(gx, gy)


def boundary(key) -> np.ndarray:
    e = np.zeros(key.shape, bool)
    d = key[:, 1:] != key[:, :-1]; e[:, 1:] |= d; e[:, :-1] |= d
    d = key[1:, :] != key[:-1, :]; e[1:, :] |= d; e[:-1, :] |= d
    return e


def align_scores(g, e, R=4, region=None) -> np.ndarray:
    """s[dy + R, dx + R] = mean reference gradient over our boundary pixels displaced by (dy, dx)"""
    m = e if region is None else (e & region)
    ys, xs = np.nonzero(m)
    H, W = g.shape
    ok = (ys >= R) & (ys < H - R) & (xs >= R) & (xs < W - R)
    ys, xs = ys[ok], xs[ok]
    s = np.zeros((2 * R + 1, 2 * R + 1))
    for dy in range(-R, R + 1):
        for dx in range(-R, R + 1):
            s[dy + R, dx + R] = g[ys + dy, xs + dx].mean()
    return s


def quadrants(shape):
    H, W = shape
    yy, xx = np.mgrid[0:H, 0:W]
    return {"all": None, "TL": (yy < H // 2) & (xx < W // 2), "TR": (yy < H // 2) & (xx >= W // 2),
            "BL": (yy >= H // 2) & (xx < W // 2), "BR": (yy >= H // 2) & (xx >= W // 2)}


def silhouette_report(ref_img, key, R=4) -> dict:
    """per region: (dy, dx) of the best alignment and the peak's ratio to the mean score 2 px away"""
    g = grad_mag(ref_img); e = boundary(key)
    out = {}
    for name, reg in quadrants(key.shape).items():
        s = align_scores(g, e, R, reg)
        j = np.unravel_index(int(np.argmax(s)), s.shape)
        out[name] = {"shift": (int(j[0]) - R, int(j[1]) - R),
                     "sharpness": float(s[R, R] / np.mean([s[R, R - 2], s[R, R + 2], s[R - 2, R], s[R + 2, R]]))}
    return out


def display_srgb8(radiance, exposure=1.0, tonemap="linear") -> np.ndarray:
    """film mean -> what the reference writes to a .png, in [0, 255] floats (not rounded)"""
    x = 1.0 - np.exp(-radiance[..., :3].astype(np.float64) * exposure)
    if tonemap == "aces":
        x = x * (2.51 * x + 0.03) / (x * (2.43 * x + 0.59) + 0.14)
    x = np.clip(x, 0.0, 1.0)
    s = np.where(x <= 0.0031308, 12.92 * x, 1.055 * np.power(np.maximum(x, 1e-12), 1 / 2.4) - 0.055)
    return 255.0 * np.clip(s, 0.0, 1.0)


def block_means(a, n) -> np.ndarray:
    h, w, c = a.shape
    return a.reshape(n, h // n, n, w // n, c).mean((1, 3))


def region_means(img, ids, which) -> dict:
    """mean colour of img over the pixels whose id (an id map at ANY resolution dividing img's) is in `which`"""
    fy, fx = img.shape[0] // ids.shape[0], img.shape[1] // ids.shape[1]
    if fy > 1 or fx > 1:
        img = img.reshape(ids.shape[0], fy, ids.shape[1], fx, -1).mean((1, 3))
    return {i: img[ids == i].mean(0) for i in which}
