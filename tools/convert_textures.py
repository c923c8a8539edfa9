"""Decode LDR/HDR textures next to a scene into side-car .pfm files for the C++ host layer
(libvision_host has no image-decoding library; image I/O is a SURVEY.md §8f "next" row).
    python tools/convert_textures.py scenes/cbox
The .pfm holds decoded values / 255 for 8-bit sources (no colour-space conversion: the host applies it)."""
import sys
from pathlib import Path

import numpy as np
from PIL import Image


def write_pfm(path, rgb):
    h, w, _ = rgb.shape
    with open(path, "wb") as f:
        f.write(f"PF\n{w} {h}\n-1.0\n".encode())
        f.write(np.ascontiguousarray(rgb[::-1].astype("<f4")).tobytes())


for d in sys.argv[1:]:
    for p in Path(d).rglob("*"):
        if p.suffix.lower() in (".jpg", ".jpeg", ".png", ".bmp", ".tga"):
            im = np.asarray(Image.open(p).convert("RGB"), np.float32) / np.float32(255.0)
            write_pfm(str(p) + ".pfm", im)
            print("wrote", str(p) + ".pfm", im.shape)
