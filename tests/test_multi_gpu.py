"""Several GPUs through the C ABI (skipped on a one-GPU box): the collective lives in libb200pt.so (ncclReduce inside
b200pt_render_frame[_all]) and is reachable from the C++ host (`path_tracer --gpus N`), as the reference's OWL context can
span devices (extern/owl/owl/RayGen.cpp:140-146; it asks for one, application.cpp:62)."""
import subprocess

import numpy as np
import pytest

import owl_path_tracer_b200 as pt
from checkers import REPO, Oracle

pytestmark = pytest.mark.gpu


def n_gpus():
    import torch
    return torch.cuda.device_count() if torch.cuda.is_available() else 0


needs2 = pytest.mark.skipif(n_gpus() < 2, reason="needs two GPUs")


@needs2
def test_two_ranks_in_one_process_equal_the_oracles_two_rank_frame():
    sc = pt.Scene(scene="cornell-box", width=96, height=64, spp=8, depth=6, env="color", env_intensity=0.0)
    ctxs = [pt.Context(device=r, rank=r, world_size=2, shard_mode=pt.SHARD_SPP) for r in range(2)]
    for c in ctxs:
        c.bind(sc)
    pt.Context.comm_init_all(ctxs)
    pt.Context.render_frame_all(ctxs)
    got = ctxs[0].framebuffer_f32()[..., :3]
    rgba = ctxs[0].framebuffer_rgba8()
    o = Oracle(det=True).load_scene(sc)
    parts = [o.render(rank=r, world=2, shard_mode=pt.SHARD_SPP) for r in range(2)]
    total = parts[0]["sum"] + parts[1]["sum"]                       # a two-rank sum has one order
    want = (total[..., :3] * (np.float32(1.0) / np.float32(8)))[::-1]
    assert (np.ascontiguousarray(got).view(np.uint32) == np.ascontiguousarray(want).view(np.uint32)).all()
    assert sum(c.stats()["rays"] for c in ctxs) == sum(p["rays"] for p in parts)
    # the pipeline over two ranks: slot j of both ranks shares a communicator, reduces run in frame order
    more = [pt.Context(device=r, rank=r, world_size=2, shard_mode=pt.SHARD_SPP) for r in range(2)]
    for c in more:
        c.bind(sc)
    pt.Context.comm_init_all(more)
    pipes = [pt.Pipeline([ctxs[r], more[r]]) for r in range(2)]
    outs = [np.zeros((64, 96), np.uint32) for _ in range(3)]
    import threading
    def drive(r):
        for k in range(3):
            pipes[r].submit(sc.materials(), outs[k] if r == 0 else None)
        pipes[r].wait()
    th = [threading.Thread(target=drive, args=(r,)) for r in range(2)]
    for t in th: t.start()
    for t in th: t.join()
    for k in range(3):
        assert (outs[k] == rgba).all(), k
    for p in pipes:
        p.close()
    for c in ctxs + more:
        c.close()


@needs2
def test_path_tracer_gpus_2_writes_the_frame_of_the_python_two_rank_path(tmp_path):
    from test_host import png_pixels
    exe = REPO / "owl_path_tracer_b200" / "bin" / "path_tracer"
    args = ["--assets", str(REPO / "assets"), "--scene", "cornell-box", "--size", "96", "64", "--spp", "8", "--depth", "6",
            "--env", "color", "--env-intensity", "0", "--out", str(tmp_path)]
    r = subprocess.run([str(exe), "--gpus", "2", *args], capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stderr[-800:]
    pngs = sorted(tmp_path.glob("*.png"))
    assert len(pngs) == 5                                            # settings.json test block: step 0.25 -> 5 frames
    sc = pt.Scene(scene="cornell-box", width=96, height=64, spp=8, depth=6, env="color", env_intensity=0.0)
    ctxs = [pt.Context(device=r, rank=r, world_size=2, shard_mode=pt.SHARD_SPP) for r in range(2)]
    for c in ctxs:
        c.bind(sc)
    pt.Context.comm_init_all(ctxs)
    pt.Context.render_frame_all(ctxs)
    want = ctxs[0].framebuffer_rgba8()
    for p in pngs:                                                   # the swept attribute (Light.subsurface) does not change the image
        assert (png_pixels(p) == want).all(), p.name
    one = tmp_path / "one"
    one.mkdir()
    r1 = subprocess.run([str(exe), *args[:-1], str(one)], capture_output=True, text=True, timeout=300)
    assert r1.returncode == 0
    assert not (png_pixels(sorted(one.glob("*.png"))[0]) == want).all()   # other RNG streams: a different (equally valid) frame
    for c in ctxs:
        c.close()
