"""Stand-in for the stripped `dragon.obj` of the reference's res/cbox-dragon.json.

The Stanford dragon (~870 k triangles) is not in the reference tree; this writes a closed, seed-free
procedural mesh of the same triangle count class: an ellipsoid (radii 0.5 / 0.33 / 0.26, the dragon's
extent in model space) on a W x H latitude-longitude grid, displaced along the radius by a fixed sum of
sinusoids (bumps from 1/3 down to 1/60 of the body), poles closed with fans. Counter-clockwise faces
seen from outside, positions only (the loader derives normals), file order = prim_id order.

  python tools/make_dragon_standin.py                 # scenes/cbox_dragon/dragon.obj        (W = 660: 869 880 triangles)
  python tools/make_dragon_standin.py --width 48      # scenes/cbox_dragon/dragon_small.obj  (W = 48:    4 512 triangles)
`convert()` is what __graft_entry__.build() calls: writes whichever of the two files is missing.
"""
from __future__ import annotations

import argparse
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent.parent
OUT_DIR = ROOT / "scenes" / "cbox_dragon"


def standin_mesh(width: int):
    """returns (positions float64 [nv,3], triangles int64 [nt,3]); nt = 2 * width * (height - 2), height = width + 1"""
    W, H = width, width + 1                      # H latitude rows including the two poles
    theta = np.pi * np.arange(1, H - 1) / (H - 1)             # interior rows
    phi = 2.0 * np.pi * np.arange(W) / W
    T, P = np.meshgrid(theta, phi, indexing="ij")

    def radius_scale(t, p):
        d = (0.10 * np.sin(3 * p + 0.5) * np.sin(2 * t) + 0.06 * np.sin(7 * p) * np.sin(5 * t + 1.0)
             + 0.03 * np.sin(17 * p + 2.0) * np.sin(13 * t) + 0.012 * np.sin(41 * p) * np.sin(37 * t + 0.3)
             + 0.004 * np.sin(97 * p + 1.0) * np.sin(89 * t))
        return 1.0 + d * np.sin(t) ** 2          # no displacement at the poles: the fans stay regular

    def point(t, p):
        s = radius_scale(t, p)
        return np.stack([0.5 * s * np.sin(t) * np.cos(p), 0.33 * s * np.cos(t), 0.26 * s * np.sin(t) * np.sin(p)], axis=-1)

    body = point(T, P).reshape(-1, 3)
    north, south = np.array([[0.0, 0.33, 0.0]]), np.array([[0.0, -0.33, 0.0]])
    pos = np.concatenate([north, body, south])
    n_rows = H - 2
    idx = 1 + np.arange(n_rows * W).reshape(n_rows, W)
    nxt = np.roll(idx, -1, axis=1)
    tris = []
    # north fan (row 0), quads between rows, south fan
    tris.append(np.stack([np.zeros(W, np.int64), nxt[0], idx[0]], axis=1))
    a, b, c, d = idx[:-1], nxt[:-1], idx[1:], nxt[1:]
    tris.append(np.stack([a, b, c], axis=-1).reshape(-1, 3))
    tris.append(np.stack([b, d, c], axis=-1).reshape(-1, 3))
    sp = len(pos) - 1
    tris.append(np.stack([np.full(W, sp, np.int64), idx[-1], nxt[-1]], axis=1))
    return pos, np.concatenate(tris).astype(np.int64)


def write_obj(path: Path, width: int) -> int:
    pos, tri = standin_mesh(width)
    path.parent.mkdir(parents=True, exist_ok=True)
    with open(path, "w") as f:
        f.write(f"# dragon stand-in, tools/make_dragon_standin.py --width {width}: {len(pos)} vertices, {len(tri)} triangles\n")
        np.savetxt(f, pos, fmt="v %.7f %.7f %.7f")
        np.savetxt(f, tri + 1, fmt="f %d %d %d")
    return len(tri)


def convert() -> None:
    for name, width in (("dragon.obj", 660), ("dragon_small.obj", 48)):
        if not (OUT_DIR / name).exists():
            write_obj(OUT_DIR / name, width)


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--width", type=int, default=0)
    ap.add_argument("--out", default="")
    a = ap.parse_args()
    if a.width:
        out = Path(a.out) if a.out else OUT_DIR / ("dragon.obj" if a.width == 660 else "dragon_small.obj")
        print(out, write_obj(out, a.width), "triangles")
    else:
        convert()
        print("wrote", sorted(p.name for p in OUT_DIR.glob("*.obj")))
