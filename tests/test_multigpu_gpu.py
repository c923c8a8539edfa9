"""GPU, >= 2 devices: the product's multi-GPU film path. Frames are partitioned by index (rank g renders
g, g + G, ...) into SUM films; the film is made whole by vb_film_reduce_peers (one process, peer memory over NVLink, fused with
the resolve) or vb_film_allreduce (ncclAllReduce). Both must equal the single-context SUM film of the same frame indices:
relMSE <= 1e-9 (fp32 sums of the same per-frame values in a different order), and the peer path must be deterministic."""
import subprocess
import threading
from pathlib import Path

import numpy as np
import pytest

from vision_b200 import Context, capi, host_api, scene_py

pytestmark = pytest.mark.gpu
ROOT = Path(__file__).resolve().parent.parent
SCENE = ROOT / "scenes" / "cbox" / "cbox_materials.json"
FRAMES = 16


def n_devices():
    import torch
    return torch.cuda.device_count()


needs2 = pytest.mark.skipif("n_devices() < 2", reason="needs 2 GPUs (gpurun --gpus 2)")


def rel_mse(a, b):
    return float(np.mean((a - b) ** 2 / (b ** 2 + 1e-2)))


def single_sum_film(sc):
    ctx = Context(0)
    ctx.upload_scene(sc)
    ctx.film_clear()
    ctx.render_frames(0, FRAMES, 1, capi.VB_ACCUM_SUM)
    acc = ctx.film_download(capi.VB_FILM_ACCUM)
    ctx.film_resolve(capi.VB_ACCUM_SUM, sc.exposure, sc.tonemap)
    out = ctx.film_download(capi.VB_FILM_OUTPUT)
    ctx.close()
    return acc, out


@needs2
def test_peer_reduce_and_nccl_allreduce_equal_single_context():
    sc = scene_py.build_scene(SCENE, {"resolution": [192, 160]})
    want_acc, want_out = single_sum_film(sc)
    assert np.all(want_acc[..., 3] == FRAMES)
    G = 2
    ctxs = [Context(g) for g in range(G)]
    for g, c in enumerate(ctxs):
        c.upload_scene(sc)
        c.film_clear()
        c.render_frames(g, FRAMES // G, G, capi.VB_ACCUM_SUM)
    parts = [c.film_download(capi.VB_FILM_ACCUM) for c in ctxs]
    assert np.all(parts[0][..., 3] == FRAMES // G)
    # (a) peer-memory fused reduce + resolve into context 0
    capi.film_reduce_peers(ctxs, 0, sc.exposure, sc.tonemap)
    acc = ctxs[0].film_download(capi.VB_FILM_ACCUM)
    out = ctxs[0].film_download(capi.VB_FILM_OUTPUT)
    assert np.array_equal(acc, parts[0] + parts[1])            # rank-order fp32 sum: exactly this
    assert np.all(acc[..., 3] == FRAMES)
    assert rel_mse(acc[..., :3], want_acc[..., :3]) <= 1e-9
    assert np.abs(out - want_out).max() <= 2e-6
    assert np.array_equal(ctxs[1].film_download(capi.VB_FILM_ACCUM), parts[1])     # the other rank's film is only read
    # (b) ncclAllReduce through the library's run-time NCCL binding, one host thread per rank
    for g, c in enumerate(ctxs):
        c.film_clear()
        c.render_frames(g, FRAMES // G, G, capi.VB_ACCUM_SUM)
    uid = capi.nccl_unique_id()
    comms, errs = [None] * G, []

    def rank_main(g):
        try:
            comms[g] = ctxs[g].nccl_comm_create(uid, G, g)
            ctxs[g].film_allreduce(comms[g])
            ctxs[g].synchronize()
        except Exception as e:      # noqa: BLE001
            errs.append(e)
    th = [threading.Thread(target=rank_main, args=(g,)) for g in range(G)]
    [t.start() for t in th]
    [t.join(timeout=120) for t in th]
    assert not errs, errs
    for c in ctxs:
        a = c.film_download(capi.VB_FILM_ACCUM)
        assert np.all(a[..., 3] == FRAMES)
        assert rel_mse(a[..., :3], want_acc[..., :3]) <= 1e-9
    for g, c in enumerate(ctxs):
        c.nccl_comm_destroy(comms[g])
        c.close()


@needs2
@pytest.mark.parametrize("nccl", [False, True], ids=["peers", "nccl"])
def test_vision_cli_two_gpus_equals_one(tmp_path, nccl):
    """the C++ host's own multi-GPU path (vh::render_on_gpus): same picture as one GPU"""
    cli = str(ROOT / "vision_b200" / "vision-cli")
    one, two = tmp_path / "one.pfm", tmp_path / "two.pfm"
    r = subprocess.run([cli, "-s", str(SCENE), "-o", str(tmp_path / "one.png"), "--pfm", str(one), "--spp", str(FRAMES)],
                       capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stderr[-2000:]
    r = subprocess.run([cli, "-s", str(SCENE), "-o", str(tmp_path / "two.png"), "--pfm", str(two), "--spp", str(FRAMES), "--gpus", "2"]
                       + (["--nccl"] if nccl else []), capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stderr[-2000:]
    assert "2 GPUs" in r.stderr
    a, b = scene_py.read_pfm(one), scene_py.read_pfm(two)
    assert a.shape == b.shape and a[..., :3].mean() > 0.01
    assert rel_mse(b[..., :3], a[..., :3]) <= 1e-9


@needs2
def test_host_pipeline_partition_with_nccl():
    """one HostPipeline per rank (what bench.py does per process), here two ranks on two host threads"""
    G = 2
    uid = capi.nccl_unique_id()
    pics, errs = [None] * G, []

    def rank_main(g):
        try:
            hp = host_api.HostPipeline(SCENE)
            hp.set_partition(g, G)
            hp.prepare(g)
            hp.nccl_init(uid)
            hp.render(FRAMES)
            pics[g] = hp.final_picture()
            hp.close()
        except Exception as e:      # noqa: BLE001
            errs.append(e)
    th = [threading.Thread(target=rank_main, args=(g,)) for g in range(G)]
    [t.start() for t in th]
    [t.join(timeout=300) for t in th]
    assert not errs, errs
    hp = host_api.HostPipeline(SCENE)
    hp.prepare(0)
    hp.render(FRAMES)
    want = hp.final_picture()
    hp.close()
    for p in pics:
        assert np.abs(p - want).max() <= 2e-6
