"""world_size-2 gloo test of the multi-GPU host logic on CPU: strided frame partition + film SUM
all-reduce reproduces the single-process film (oracle stands in for the device renderer here)."""
import os
import sys
from pathlib import Path

import numpy as np
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT)); sys.path.insert(0, str(ROOT / "tests"))

from vision_b200.dist import frames_for_rank, ray_range_for_rank


def test_partitions_cover_exactly():
    for world in (1, 2, 3, 4, 8):
        for spp in (1, 2, 7, 16, 1024):
            seen = []
            for r in range(world):
                f, n, s = frames_for_rank(r, world, spp)
                seen += [f + i * s for i in range(n)]
            assert sorted(seen) == list(range(spp))
        for n_rays in (0, 1, 10, 1 << 20):
            cuts = [ray_range_for_rank(r, world, n_rays) for r in range(world)]
            assert cuts[0][0] == 0 and cuts[-1][1] == n_rays
            assert all(cuts[i][1] == cuts[i + 1][0] for i in range(world - 1))


def _worker(rank, world, port, out_path):
    os.environ["MASTER_ADDR"] = "127.0.0.1"; os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    import oracle_lib
    from vision_b200 import scene_py
    sc = scene_py.build_scene(ROOT / "scenes" / "cbox" / "cbox_matte.json", dict(scene_py.CBOX_MATTE_OVERRIDES, resolution=[32, 32], spp=8))
    osc = oracle_lib.OracleScene(oracle_lib.load(), sc, threads=2)
    first, n, stride = frames_for_rank(rank, world, sc.spp)
    film, _ = osc.render(first, n, stride=stride, mode=1, threads=2)          # SUM film of this rank's frames
    t = torch.from_numpy(film)
    dist.all_reduce(t)                                                          # the one exchange step of the path
    if rank == 0:
        np.save(out_path, t.numpy())
    dist.destroy_process_group()


def test_two_rank_film_allreduce_matches_single(tmp_path):
    out = tmp_path / "film.npy"
    mp.spawn(_worker, args=(2, 29533, str(out)), nprocs=2, join=True)
    summed = np.load(out)
    import oracle_lib
    from vision_b200 import scene_py
    sc = scene_py.build_scene(ROOT / "scenes" / "cbox" / "cbox_matte.json", dict(scene_py.CBOX_MATTE_OVERRIDES, resolution=[32, 32], spp=8))
    osc = oracle_lib.OracleScene(oracle_lib.load(), sc, threads=2)
    ref, _ = osc.render(0, 8, mode=0, threads=2)                               # reference running-mean film
    assert np.allclose(summed[..., 3], 8.0)
    mean = summed[..., :3] / summed[..., 3:4]
    err = float(np.mean((mean - ref[..., :3]) ** 2 / (ref[..., :3] ** 2 + 1e-2)))
    assert err < 1e-10, err
