"""Evaluates the reference's measured complex IOR tables (numeric data in
src/render_core/material/metal_ior.inl.h) at the three sRGB peak wavelengths, the way MetalMaterial
does for a non-spectral ("srgb") spectrum: SPD::eval (src/base/color/spd.cpp:67-76) at
rgb_spectrum_peak_wavelengths (src/base/color/spectrum.h:14). Writes vision_b200/data/metal_ior_srgb.json.
Run once in the build container. Usage: python tools/extract_metal_ior.py [/root/reference]"""
import json
import re
import sys
from pathlib import Path

import numpy as np

f32 = np.float32
ref = Path(sys.argv[1] if len(sys.argv) > 1 else "/root/reference")
src = (ref / "src/render_core/material/metal_ior.inl.h").read_text()
PEAKS = (602.785, 539.285, 445.772)
LMIN, LMAX, CIE_N = f32(360.0), f32(830.0), 471


def spd_eval(func, lam):
    interval = f32(CIE_N) / f32(len(func))
    t = (min(max(f32(lam), LMIN), LMAX) - LMIN) / interval
    sample_count = f32(int((LMAX - LMIN) / interval) + 1)
    i = int(min(f32(t), f32(sample_count - f32(2))))
    l, r = f32(func[i]), f32(func[i + 1])
    fr = f32(t) - f32(np.floor(f32(t)))
    return float(f32(l + fr * (r - l)))


out = {}
for m in re.finditer(r"ocarina::array\s+(\w+)\s*\{(.*?)\};", src, re.S):
    pairs = re.findall(r"make_float2\(\s*([-+0-9.eE]+)f?\s*,\s*([-+0-9.eE]+)f?\s*\)", m.group(2))
    eta = [float(f32(a)) for a, _ in pairs]
    k = [float(f32(b)) for _, b in pairs]
    out[m.group(1)] = {"eta": [spd_eval(eta, l) for l in PEAKS], "k": [spd_eval(k, l) for l in PEAKS], "samples": len(pairs)}
dst = Path(__file__).resolve().parent.parent / "vision_b200" / "data" / "metal_ior_srgb.json"
dst.write_text(json.dumps(out, indent=1))
# the full tables, for the "spd" eta / k slots MetalMaterial creates under a complete (hero) spectrum (metal.cpp:117-121)
full = {}
for m in re.finditer(r"ocarina::array\s+(\w+)\s*\{(.*?)\};", src, re.S):
    pairs = re.findall(r"make_float2\(\s*([-+0-9.eE]+)f?\s*,\s*([-+0-9.eE]+)f?\s*\)", m.group(2))
    full[m.group(1)] = {"eta": [float(f32(a)) for a, _ in pairs], "k": [float(f32(b)) for _, b in pairs]}
(dst.parent / "metal_ior_spd.json").write_text(json.dumps(full))
for k_, v in out.items():
    print(k_, v)
