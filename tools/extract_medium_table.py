"""Converts the measured subsurface parameter table (numeric data in
src/base/import/medium_scatter_data.h, used by MediumDesc::init for "medium_name") into
vision_b200/data/medium_scatter.json: {name: {"sigma_s": [3], "sigma_a": [3]}} in table order.
Run once in the build container. Usage: python tools/extract_medium_table.py [/root/reference]"""
import json
import re
import sys
from pathlib import Path

ref = Path(sys.argv[1] if len(sys.argv) > 1 else "/root/reference")
src = (ref / "src/base/import/medium_scatter_data.h").read_text()
num = r"([-+0-9.eE]+)"
out = {}
for m in re.finditer(r'\{"([^"]+)",\s*float3\(' + num + r",\s*" + num + r",\s*" + num + r"\),\s*float3\(" + num + r",\s*" + num + r",\s*" + num + r"\)\}", src):
    out[m.group(1)] = {"sigma_s": [float(m.group(i)) for i in (2, 3, 4)], "sigma_a": [float(m.group(i)) for i in (5, 6, 7)]}
dst = Path(__file__).resolve().parent.parent / "vision_b200" / "data" / "medium_scatter.json"
dst.write_text(json.dumps(out, indent=1))
print(len(out), "entries; first:", next(iter(out)))
