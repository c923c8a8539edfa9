"""Converts the CIE 1931 observer and D65 illuminant tables (numeric data, 1 nm steps 360..830 nm, in
src/base/color/cie.h:22-412) into vision_b200/data/cie_1nm_4x471.f32: rows X, Y, Z, D65, float32.
Run once in the build container. Usage: python tools/extract_cie.py [/root/reference]"""
import re
import sys
from pathlib import Path

import numpy as np

ref = Path(sys.argv[1] if len(sys.argv) > 1 else "/root/reference")
src = (ref / "src/base/color/cie.h").read_text()
rows = []
for name in ("X", "Y", "Z", "D65"):
    m = re.search(r"array<float,\s*cie_sample_count>\s+" + name + r"\s*=\s*\{(.*?)\};", src, re.S)
    vals = [float(v.rstrip("f")) for v in re.findall(r"[-+]?[0-9]*\.?[0-9]+(?:[eE][-+]?[0-9]+)?f?", m.group(1))]
    assert len(vals) == 471, (name, len(vals))
    rows.append(vals)
dst = Path(__file__).resolve().parent.parent / "vision_b200" / "data" / "cie_1nm_4x471.f32"
np.asarray(rows, dtype=np.float32).tofile(dst)
print("wrote", dst, "Y max", max(rows[1]), "D65[200]", rows[3][200])
