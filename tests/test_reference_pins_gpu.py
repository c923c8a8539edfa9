"""GPU: the CUDA path, loaded and driven by the C++ HOST LAYER (libvision_host / vision-cli, not the Python twin), against
  (1) images the reference itself rendered (tests/golden/reference_renders/): silhouettes and radiometry, same measurements and
      tolerances as tests/test_reference_pins.py applies to the oracle;
  (2) the oracle fed with the records the C++ host produced, at the BASELINE configs' FULL resolution
      (stock cbox 1024 x 1024, classroom 1920 x 1080): primary-hit ids bit-exact over every pixel, films bit-identical (classroom: all 2 073 600 pixels of a 2-spp film)."""
import subprocess
from pathlib import Path

import numpy as np
import pytest

import refpin
from test_reference_pins import check_radiometry, check_silhouettes
from vision_b200 import capi, host_api

pytestmark = pytest.mark.gpu
ROOT = Path(__file__).resolve().parent.parent


def host_pipeline(path, overrides=None):
    hp = host_api.HostPipeline(ROOT / path, overrides)
    hp.prepare(0)
    return hp


def ids_of(hits, w, h):
    return hits["inst"].reshape(h, w).astype(np.int64), hits["prim"].reshape(h, w).astype(np.int64)


def test_cbox_stock_full_size_vs_reference_render_and_oracle(oracle):
    import oracle_lib
    hp = host_pipeline("scenes/cbox_stock/vision_scene.json")                 # the reference's stock file: 1024 x 1024
    i = hp.info()
    assert (i["width"], i["height"]) == (1024, 1024)
    g = hp.context().dump_primary_hits(0)
    inst, prim = ids_of(g, 1024, 1024)
    ref = refpin.load_png(refpin.GOLDEN / "cbox_srgb.png")
    rep = refpin.silhouette_report(ref, inst * 16 + prim // 2)
    print(rep)
    check_silhouettes(rep, "cbox")
    # the same frame from the oracle, over the records the C++ host built: every pixel's ids and hit values identical
    osc = oracle_lib.OracleScene(oracle, hp.records(), threads=oracle.hw_threads())
    o = osc.primary_hits(0, threads=oracle.hw_threads())
    for f in ("inst", "prim", "u", "v", "t"):
        assert g[f].tobytes() == o[f].tobytes(), f
    hp.close()


def test_cbox_lookalike_through_vision_cli_matches_reference_render(tmp_path, oracle):
    """the whole product path a user runs: vision-cli -> scene import -> CUDA render -> exposure / tone map -> gamma -> .png"""
    out = tmp_path / "lookalike.png"
    spp = 256
    r = subprocess.run([str(ROOT / "vision_b200" / "vision-cli"), "-s", str(ROOT / "scenes/cbox_stock/lookalike.json"),
                        "-o", str(out), "--spp", str(spp)], capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stderr[-2000:]
    img8 = refpin.load_png(out)
    ref = refpin.load_png(refpin.GOLDEN / "cbox_srgb.png")
    assert img8.shape == ref.shape == (1024, 1024, 3)
    hp = host_pipeline("scenes/cbox_stock/lookalike.json")
    inst, _ = ids_of(hp.context().dump_primary_hits(0), 1024, 1024)
    hp.close()
    check_radiometry(img8, inst, ref)
    # full-resolution figure for the record (both images are Monte Carlo estimates at different sample counts and seeds)
    a, b = refpin.block_means(img8, 256), refpin.block_means(ref, 256)
    print(f"256x256 block means (4x4 px): mean |diff| {np.abs(a - b).mean():.2f} / 255, "
          f"relMSE {np.mean((a - b) ** 2 / (b ** 2 + 25.0)):.4f}")
    assert np.abs(a - b).mean() <= 8.0


def test_classroom_full_size_vs_reference_render_and_oracle(oracle):
    """config C5 as stated (1920 x 1080, depth 16, fog + environment), loaded by the C++ host: one frame vs the oracle over ALL
    pixels; and the silhouettes of the 1280 x 720 framing the reference's own render used"""
    import oracle_lib
    ref = refpin.load_png(refpin.GOLDEN / "classroom_1024spp.png")
    hp = host_pipeline("scenes/classroom/vision_scene.json", {"resolution": [1280, 720]})
    inst, _ = ids_of(hp.context().dump_primary_hits(0), 1280, 720)
    hp.close()
    rep = refpin.silhouette_report(ref, inst)
    print(rep)
    check_silhouettes(rep, "classroom")

    hp = host_pipeline("scenes/classroom/vision_scene.json", {"resolution": [1920, 1080]})
    ctx = hp.context()
    g = ctx.dump_primary_hits(0)
    hp.render(2)                                    # frames 0, 1 folded by the reference's running lerp (frame_buffer.cpp:117-126)
    film = hp.accum()
    stats = ctx.render_stats()
    osc = oracle_lib.OracleScene(oracle, hp.records(), threads=oracle.hw_threads())
    o = osc.primary_hits(0, threads=oracle.hw_threads())
    for f in ("inst", "prim", "u", "v", "t"):
        assert g[f].tobytes() == o[f].tobytes(), f
    ofilm, _, ost = osc.render(0, 2, threads=oracle.hw_threads(), stats=True)
    same = (film == ofilm).all(axis=-1).mean()
    err = float(np.mean((film[..., :3] - ofilm[..., :3]) ** 2 / (ofilm[..., :3] ** 2 + 1e-2)))
    print(f"classroom 1920x1080 x 2 spp: bit-identical pixels {same:.6f}, relMSE {err:.3e}, closest rays {stats['closest']} vs {ost['closest']}")
    assert same == 1.0 and err == 0.0, (same, err)       # every one of the 2 073 600 pixels, bit for bit
    # the reference (and the oracle) trace the camera ray twice, in rt_geom and again as bounce 0 of Li; the CUDA path traces it once
    assert stats["closest"] + 2 * 1920 * 1080 == ost["closest"]
    hp.close()
