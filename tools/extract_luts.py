"""Converts the energy-compensation tables of the reference (numeric data in
src/base/scattering/precomputed_table.h) into raw little-endian float32 files under
vision_b200/data/. Run once in the build container (the reference tree is not on the GPU box);
the .f32 files are committed as data. Usage: python tools/extract_luts.py [/root/reference]"""
import re
import sys
from pathlib import Path

import numpy as np

ref = Path(sys.argv[1] if len(sys.argv) > 1 else "/root/reference")
src = (ref / "src/base/scattering/precomputed_table.h").read_text()
out = Path(__file__).resolve().parent.parent / "vision_b200" / "data"
out.mkdir(exist_ok=True)
want = {"PureReflectionLobe_Table": ("lut_pure_reflection_32x32x1x1.f32", 1024, 1),
        "DielectricLobe_Table": ("lut_dielectric_32x32x32x2.f32", 32768, 2),
        "DielectricInvLobe_Table": ("lut_dielectric_inv_32x32x32x2.f32", 32768, 2),
        "SpecularLobe_Table": ("lut_specular_32x32x32x1.f32", 32768, 1),
        "CoatLobe_Table": ("lut_coat_32x32x32x1.f32", 32768, 1),
        # principled sheen (Zeltner et al. 2022 LTC fits), src/render_core/material/ltc_sheen_table.h
        "SheenLTCTableVolume": ("lut_sheen_volume_32x32x1x4.f32", 1024, 4),
        "SheenLTCTableApprox": ("lut_sheen_approx_32x32x1x4.f32", 1024, 4)}
src += (ref / "src/render_core/material/ltc_sheen_table.h").read_text()
for name, (fn, count, ch) in want.items():
    m = re.search(r"const\s+\w+\s+" + name + r"\[(\d+)\]\s*=\s*\{(.*?)\};", src, re.S)
    assert m and int(m.group(1)) == count, name
    nums = re.findall(r"[-+]?(?:\d+\.\d*|\.\d+|\d+)(?:[eE][-+]?\d+)?", re.sub(r"(\d)f\b", r"\1", re.sub(r"float\d", "", m.group(2))))
    a = np.array([float(x) for x in nums], np.float32)
    assert a.size == count * ch, (name, a.size)
    a.tofile(out / fn)
    print(fn, a.size, float(a.min()), float(a.max()))
