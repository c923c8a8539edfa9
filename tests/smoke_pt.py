"""second half of __graft_entry__.smoke(): one tiny path-traced frame batch vs the oracle."""
from pathlib import Path

import numpy as np


def run(ctx, oracle):
    import oracle_lib
    from vision_b200 import scene_py, capi
    root = Path(__file__).resolve().parent.parent
    sc = scene_py.build_scene(root / "scenes" / "cbox" / "cbox_matte.json",
                              dict(scene_py.CBOX_MATTE_OVERRIDES, resolution=[64, 64], spp=2))
    ctx.upload_scene(sc)
    ctx.render_frames(0, 2, 1, capi.VB_ACCUM_LERP)
    g = ctx.film_download(capi.VB_FILM_ACCUM)
    osc = oracle_lib.OracleScene(oracle, sc, threads=4)
    o, _ = osc.render(0, 2, threads=4)
    err = float(np.mean((g[..., :3] - o[..., :3]) ** 2 / (o[..., :3] ** 2 + 1e-2)))
    assert err <= 1e-6, f"smoke: path-traced image differs from oracle (relMSE {err})"
    print(f"smoke ok: cbox matte 64x64 x 2 spp, relMSE vs oracle {err:.2e}, mean radiance {g[..., :3].mean():.4f}")
