"""CPU: the ORACLE against images the reference itself rendered (tests/golden/reference_renders/, copied byte for byte by
tools/import_reference_renders.py). These are the only reference-produced outputs that exist offline; they
pin the conventions SURVEY.md Appendix A1 could only infer:
  silhouettes  -> look_at / perspective / scale(1,1,-1) handedness, matrix4x4 column order, fov_y, raster orientation, quad and
                  cube vertex layout, OBJ model placement (classroom: 79 models)          [sensor.cpp:58-78,153-162, node_desc.cpp:31-58]
  radiometry   -> area-light emission and sampling, diffuse albedo, multi-bounce transport energy, `1 - exp(-L * exposure)`,
                  the sRGB OETF on image write                                             [integrator.cpp:160-311, frame_buffer.cpp:72-74,146-154]
Tolerances are stated in BASELINE.md section 3 and asserted here."""
from pathlib import Path

import numpy as np
import pytest

import refpin
from vision_b200 import scene_py

ROOT = Path(__file__).resolve().parent.parent

# silhouettes: best alignment must be the zero shift everywhere; the peak must stand out against a 2-pixel displacement
SHARPNESS_MIN = {"cbox": 1.3, "classroom": 1.12}
# radiometry (8-bit sRGB units, reference image is a noisy ~1k-spp render)
WALL_TOL = {0: 5.0, 2: 5.0, 3: 5.0, 4: 5.0, 1: 9.0}      # floor, back, left, right walls; ceiling (lit by the glass box's caustic only)
BLOCK_MEAN_ABS_TOL = 6.0                                  # mean |difference| over the 64 x 64 grid of block means


def primary_ids(oracle, path, res):
    import oracle_lib
    sc = scene_py.build_scene(ROOT / path, {"resolution": list(res), "spp": 1})
    osc = oracle_lib.OracleScene(oracle, sc, threads=oracle.hw_threads())
    h = osc.primary_hits(0, threads=oracle.hw_threads())
    shape = (res[1], res[0])
    return h["inst"].reshape(shape).astype(np.int64), h["prim"].reshape(shape).astype(np.int64)


def check_silhouettes(rep, name):
    for region, r in rep.items():
        assert r["shift"] == (0, 0), f"{name}/{region}: reference edges sit at {r['shift']} px from ours"
        assert r["sharpness"] >= SHARPNESS_MIN[name], f"{name}/{region}: alignment peak not distinct ({r['sharpness']:.3f})"


def test_cbox_silhouettes_match_reference_render(oracle):
    ref = refpin.load_png(refpin.GOLDEN / "cbox_srgb.png")
    inst, prim = primary_ids(oracle, "scenes/cbox_stock/vision_scene.json", (1024, 1024))
    key = inst * 16 + prim // 2                      # a cube face = two triangles
    rep = refpin.silhouette_report(ref, key)
    print(rep)
    check_silhouettes(rep, "cbox")
    # the measurement is sensitive: with our id map displaced by (+2, -2) px the reference's edges are found at (-2, +2) from it
    moved = refpin.silhouette_report(ref, np.roll(key, (2, -2), axis=(0, 1)))
    assert moved["all"]["shift"] == (-2, 2)
    # the emitter is the one region both images define exactly (saturated white): masks agree to 1 px
    light = inst == 7
    white = ref.min(-1) >= 250
    ys, xs = np.nonzero(light); yr, xr = np.nonzero(white)
    assert abs(ys.min() - yr.min()) <= 1 and abs(ys.max() - yr.max()) <= 1
    assert abs(xs.min() - xr.min()) <= 1 and abs(xs.max() - xr.max()) <= 1
    assert (light & white).sum() / (light | white).sum() > 0.98


def test_classroom_silhouettes_match_reference_render(oracle):
    ref = refpin.load_png(refpin.GOLDEN / "classroom_1024spp.png")
    inst, _ = primary_ids(oracle, "scenes/classroom/vision_scene.json", (1280, 720))      # the file's pipeline-block resolution
    rep = refpin.silhouette_report(ref, inst)
    print(rep)
    check_silhouettes(rep, "classroom")


def check_radiometry(img8, ids, ref):
    """img8: our display-referred image (float, 0..255) at a resolution dividing the reference's 1024^2"""
    a = refpin.block_means(img8, 64)
    b = refpin.block_means(ref, 64)
    mean_abs = float(np.abs(a - b).mean())
    ours = refpin.region_means(img8, ids, WALL_TOL)
    refs = refpin.region_means(ref, ids, WALL_TOL)
    for i, tol in WALL_TOL.items():
        print(f"region {i}: ours {ours[i].round(1)}  reference {refs[i].round(1)}")
    print(f"64x64 block means: mean |diff| {mean_abs:.2f} / 255")
    for i, tol in WALL_TOL.items():
        assert np.abs(ours[i] - refs[i]).max() <= tol, (i, ours[i], refs[i])
    assert mean_abs <= BLOCK_MEAN_ABS_TOL
    # and the image is not trivially flat: the reference's red / green wall contrast is there
    assert ours[3][0] > 3 * ours[3][1] and ours[4][1] > 1.3 * ours[4][0]


def test_cbox_radiometry_matches_reference_render(oracle):
    import oracle_lib
    ref = refpin.load_png(refpin.GOLDEN / "cbox_srgb.png")
    res, spp = 256, 128
    sc = scene_py.build_scene(ROOT / "scenes/cbox_stock/lookalike.json", {"resolution": [res, res], "spp": spp})
    osc = oracle_lib.OracleScene(oracle, sc, threads=oracle.hw_threads())
    film, _ = osc.render(0, spp, threads=oracle.hw_threads())
    h = osc.primary_hits(0, threads=oracle.hw_threads())
    ids = h["inst"].reshape(res, res).astype(np.int64)
    img8 = refpin.display_srgb8(film, sc.exposure, "linear")
    check_radiometry(img8, ids, ref)
    # the stock file's ACES curve is NOT what produced srgb.png: with it the walls are off by more than the tolerance
    aces8 = refpin.display_srgb8(film, sc.exposure, "aces")
    assert np.abs(refpin.region_means(aces8, ids, [2])[2] - refpin.region_means(ref, ids, [2])[2]).max() > 2 * WALL_TOL[2]
