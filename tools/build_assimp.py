"""Builds the reference's VENDORED importer library (src/ext/assimp, the code behind its `model` shape:
src/importers/assimp_parser.cpp) from where it lies in the reference tree, OBJ + PLY importers only, static, PIC, into
vision_b200/third_party/_assimp/{lib,include} (git-ignored; travels to the GPU box with the other built files). No reference or
assimp source enters the repository; libvision_host links the archive when it is present (vision_b200/host/Makefile) and the
`model` plugin then yields the reference's vertex / triangle ORDER. ~35 s on 8 cores, offline.
Run by __graft_entry__.build() when /root/reference is present and the archive is missing."""
import os
import shutil
import subprocess
import sys
import tempfile
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
OUT = ROOT / "vision_b200" / "third_party" / "_assimp"


def build(ref="/root/reference", force=False) -> bool:
    src = Path(ref) / "src" / "ext" / "assimp"
    lib = OUT / "lib" / "libassimp.a"
    if lib.exists() and not force:
        return True
    if not (src / "CMakeLists.txt").exists() or shutil.which("cmake") is None:
        return False
    tmp = Path(tempfile.mkdtemp(prefix="vb_assimp_"))
    try:
        cfg = ["cmake", str(src), "-G", "Unix Makefiles", "-DCMAKE_BUILD_TYPE=Release", "-DBUILD_SHARED_LIBS=OFF",
               "-DASSIMP_BUILD_TESTS=OFF", "-DASSIMP_BUILD_ASSIMP_TOOLS=OFF", "-DASSIMP_BUILD_SAMPLES=OFF", "-DASSIMP_NO_EXPORT=ON",
               "-DASSIMP_BUILD_ALL_IMPORTERS_BY_DEFAULT=OFF", "-DASSIMP_BUILD_OBJ_IMPORTER=ON", "-DASSIMP_BUILD_PLY_IMPORTER=ON",
               "-DASSIMP_BUILD_ZLIB=ON", "-DASSIMP_WARNINGS_AS_ERRORS=OFF", "-DASSIMP_INSTALL=OFF", "-DCMAKE_POSITION_INDEPENDENT_CODE=ON"]
        subprocess.run(cfg, cwd=tmp, check=True, capture_output=True)
        subprocess.run(["make", "-j", str(os.cpu_count() or 4)], cwd=tmp, check=True, capture_output=True)
        (OUT / "lib").mkdir(parents=True, exist_ok=True)
        shutil.copyfile(tmp / "lib" / "libassimp.a", lib)
        shutil.copyfile(tmp / "contrib" / "zlib" / "libzlibstatic.a", OUT / "lib" / "libzlibstatic.a")
        inc = OUT / "include"
        if inc.exists():
            shutil.rmtree(inc)
        shutil.copytree(src / "include", inc)                       # public headers (third-party, not committed)
        for gen in ("config.h", "revision.h"):
            shutil.copyfile(tmp / "include" / "assimp" / gen, inc / "assimp" / gen)
        (OUT / "PROVENANCE.txt").write_text(f"built from {src} by tools/build_assimp.py (OBJ + PLY importers, static, PIC)\n")
        return True
    finally:
        shutil.rmtree(tmp, ignore_errors=True)


if __name__ == "__main__":
    ok = build(sys.argv[1] if len(sys.argv) > 1 else "/root/reference", force="--force" in sys.argv)
    print("assimp:", "built" if ok else "unavailable (reference tree or cmake missing)")
