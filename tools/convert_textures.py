"""Decode LDR textures next to a scene into side-car binary .ppm (P6, 8-bit) files for the C++ host layer
(libvision_host has no image-decoding library; image I/O is a SURVEY.md §8f "next" row).
    python tools/convert_textures.py scenes/cbox scenes/classroom
The side-car holds the decoded 8-bit values (no colour-space conversion: the host divides by 255 and applies it).
__graft_entry__.build() runs this; the side-cars are git-ignored."""
import sys
from pathlib import Path

import numpy as np
from PIL import Image


def convert(dirs):
    out = []
    for d in dirs:
        for p in sorted(Path(d).rglob("*")):
            if p.suffix.lower() in (".jpg", ".jpeg", ".png", ".bmp", ".tga"):
                dst = Path(str(p) + ".ppm")
                if dst.exists() and dst.stat().st_mtime >= p.stat().st_mtime:
                    continue
                im = np.asarray(Image.open(p).convert("RGB"), np.uint8)
                with open(dst, "wb") as f:
                    f.write(f"P6\n{im.shape[1]} {im.shape[0]}\n255\n".encode())
                    f.write(np.ascontiguousarray(im).tobytes())
                out.append(str(dst))
    return out


if __name__ == "__main__":
    for f in convert(sys.argv[1:]):
        print("wrote", f)
