"""Progressive surface: b200pt_render_pass / b200pt_render_progressive / b200pt_set_output.

The reference draws all samples of a pixel from ONE RNG stream in order (device.cu:226-243), so the frame after
k samples of a progressive render must be the reference's frame at max_samples = k — bit for bit against the
deterministic-math oracle — and the last pass must leave exactly the frame of the one-shot b200pt_render.
The convergence figure (rel_variance) is checked against its numpy restatement fed with the ORACLE's first and
second moments (tolerance 1e-9 relative: binary64 sums in a different order)."""
import numpy as np
import pytest

import owl_path_tracer_b200 as pt
from checkers import Oracle, rel_variance

CORNELL = dict(scene="cornell-box", depth=8, env="color", env_intensity=0.0)


def bits(a):
    return np.ascontiguousarray(a).view(np.uint32)


# ---------------------------------------------------------------- CPU: the oracle side of the feature
def test_oracle_second_moment_and_gamma_output():
    sc = pt.Scene(width=32, height=32, spp=1, **CORNELL)
    o = Oracle(det=True).load_scene(sc)
    r1 = o.render()
    s = r1["sum"]
    lum = (np.float32(0.2126) * s[..., 0] + np.float32(0.7152) * s[..., 1]) + np.float32(0.0722) * s[..., 2]
    assert (bits(lum * lum) == bits(s[..., 3])).all()          # one sample: w = lum(value)^2, same float ops
    assert rel_variance(s, 1) == 0.0
    sc4 = pt.Scene(width=32, height=32, spp=4, **CORNELL)
    o4 = Oracle(det=True).load_scene(sc4)
    r4 = o4.render()
    assert (r4["sum"][..., 3] >= 0).all() and 0.0 < rel_variance(r4["sum"], 4) < 10.0
    # Cauchy-Schwarz: (sum of lum)^2 <= n * sum of lum^2 (up to float rounding)
    l4 = 0.2126 * r4["sum"][..., 0].astype(np.float64) + 0.7152 * r4["sum"][..., 1] + 0.0722 * r4["sum"][..., 2]
    assert (l4 * l4 <= 4.0 * r4["sum"][..., 3] * (1 + 1e-5) + 1e-12).all()
    # gamma output: the float frame is untouched, the RGBA8 frame is make_rgba(saturate(pow(mean, 1/2.2)))
    g = o4.set_output(1).render()
    assert (bits(g["mean"]) == bits(r4["mean"])).all()
    m = r4["mean"][..., :3]
    p = o4.math("pow", m.reshape(-1), np.full(m.size, np.float32(1.0) / np.float32(2.2), np.float32)).reshape(m.shape)
    p = np.minimum(np.maximum(p, np.float32(0)), np.float32(1))
    q = np.minimum(255, np.maximum(0, (p * np.float32(256)).astype(np.int32))).astype(np.uint32)
    want = q[..., 0] | (q[..., 1] << 8) | (q[..., 2] << 16) | np.uint32(0xFF000000)
    assert (g["rgba"] == want).all()
    assert (g["rgba"] != r4["rgba"]).any()


# ---------------------------------------------------------------- GPU
@pytest.mark.gpu
@pytest.mark.parametrize("flags", [0, pt.FLAG_NEE])
def test_every_pass_is_the_reference_frame_at_that_sample_count(flags):
    sc = pt.Scene(width=96, height=96, spp=8, **CORNELL)
    ctx = pt.Context(flags=flags)
    ctx.bind(sc)
    done, rays = 0, 0
    for n in (3, 3, 5):                      # the last pass is clamped to the 2 samples that are left
        info = ctx.render_pass(n)
        done = min(8, done + n)
        rays += info["rays"]
        assert info["samples_done"] == done and info["complete"] == (done == 8)
        o = Oracle(det=True).load_scene(pt.Scene(width=96, height=96, spp=done, **CORNELL)).set_nee(bool(flags))
        ro = o.render()
        assert (bits(ctx.framebuffer_f32()[..., :3]) == bits(ro["mean"][..., :3])).all(), f"after {done} samples"
        assert (ctx.framebuffer_rgba8() == ro["rgba"]).all()
        assert rays == ro["rays"]
        assert info["rel_variance"] == pytest.approx(rel_variance(ro["sum"], done), rel=1e-9)
    with pytest.raises(pt.B200ptError, match="all max_samples"):
        ctx.render_pass(1)
    frame = ctx.framebuffer_f32().copy()
    ctx.render()                              # the one-shot frame
    assert (bits(ctx.framebuffer_f32()) == bits(frame)).all()
    ctx.close()


@pytest.mark.gpu
def test_passes_on_two_concurrent_sub_frames_equal_the_one_shot_frame():
    # >= 2^18 lanes: the frame is walked as two interleaved sub-frames on two streams; walks get parked
    sc = pt.Scene(scene="dragon", width=640, height=416, spp=6, depth=8, env="color", env_intensity=0.0)
    ctx = pt.Context()
    ctx.bind(sc)
    ctx.render()
    want_f, want_q, want_rays = ctx.framebuffer_f32().copy(), ctx.framebuffer_rgba8().copy(), ctx.stats()["rays"]
    rays = 0
    for n in (1, 2, 3):
        rays += ctx.render_pass(n)["rays"]
    assert rays == want_rays
    assert (bits(ctx.framebuffer_f32()) == bits(want_f)).all() and (ctx.framebuffer_rgba8() == want_q).all()
    ctx.close()


@pytest.mark.gpu
def test_progressive_loop_early_stop_callback_and_restart():
    sc = pt.Scene(width=64, height=64, spp=64, **CORNELL)
    ctx = pt.Context()
    ctx.bind(sc)
    seen = []
    last = ctx.render_progressive(4, 0.0, lambda info, rgba, mean: seen.append((info["samples_done"], int(rgba[0, 0]), float(mean[..., :3].mean()))) and False)
    assert [s[0] for s in seen] == list(range(4, 65, 4)) and last["complete"] == 1 and last["stopped_early"] == 0
    full = ctx.framebuffer_f32().copy()
    ctx.render()
    assert (bits(full) == bits(ctx.framebuffer_f32())).all()
    # In the cornell box (tiny emitter, no light sampling) the figure RISES over the first passes: most pixels have
    # not seen the light yet and report no variance at all; it stays near the measured relMSE of such frames (0.7 at
    # 64 spp).  A well-lit scene shows the 1/n law; a target stops the loop at the first pass at or below it.
    v = []
    ctx.render_progressive(8, 0.0, lambda info, *_: v.append(info["rel_variance"]) and False)
    assert len(v) == 8 and all(x > 0 for x in v)
    cube = pt.Scene(scene="cube", width=64, height=64, spp=64, depth=8, env="auto", env_intensity=1.0)
    cc = pt.Context()
    cc.bind(cube)
    v = []
    cc.render_progressive(8, 0.0, lambda info, *_: v.append(info["rel_variance"]) and False)
    assert all(b < a for a, b in zip(v, v[1:])) and 6.0 < v[0] / v[-1] < 10.0       # 8 -> 64 samples: ~ 1/8
    target = 0.5 * (v[1] + v[2])
    early = cc.render_progressive(8, target)
    assert early["samples_done"] == 24 and early["rel_variance"] == v[2]
    assert early["stopped_early"] == 1 and early["complete"] == 0
    o = Oracle(det=True).load_scene(pt.Scene(scene="cube", width=64, height=64, spp=24, depth=8, env="auto", env_intensity=1.0))
    assert (bits(cc.framebuffer_f32()[..., :3]) == bits(o.render()["mean"][..., :3])).all()
    cc.close()
    # callback stop
    n = []
    stopped = ctx.render_progressive(8, 0.0, lambda info, *_: n.append(1) or len(n) == 2)
    assert stopped["samples_done"] == 16 and stopped["complete"] == 0
    # a scene change drops the frame in progress: the next pass starts at sample 0
    ctx.render_restart()
    ctx.render_pass(8)
    ctx.set_camera(*sc.camera)
    assert ctx.render_pass(8)["samples_done"] == 8
    # a larger sample budget extends the frame in progress
    ctx.set_render(64, 64, 128, 8)
    assert ctx.render_pass(8)["samples_done"] == 16
    ctx.close()


@pytest.mark.gpu
def test_gamma_output_and_errors():
    sc = pt.Scene(width=80, height=60, spp=4, **CORNELL)
    ctx = pt.Context()
    ctx.bind(sc)
    ctx.render()
    lin_f, lin_q = ctx.framebuffer_f32().copy(), ctx.framebuffer_rgba8().copy()
    ctx.set_output(pt.TRANSFER_GAMMA22)
    ctx.render()
    o = Oracle(det=True).load_scene(sc).set_output(1)
    ro = o.render()
    assert (bits(ctx.framebuffer_f32()) == bits(lin_f)).all()          # the float frame stays linear
    assert (ctx.framebuffer_rgba8() == ro["rgba"]).all() and (ctx.framebuffer_rgba8() != lin_q).any()
    with pytest.raises(pt.B200ptError, match="unknown transfer"):
        ctx.set_output(7)
    with pytest.raises(pt.B200ptError, match="positive"):
        ctx.render_pass(0)
    ctx.close()
    two = pt.Context(rank=0, world_size=2, shard_mode=pt.SHARD_ROWS)
    two.bind(sc)
    with pytest.raises(pt.B200ptError, match="SPP sharding"):
        two.render_pass(1)
    with pytest.raises(pt.B200ptError, match="one rank only"):
        two.render_progressive(1)
    two.close()
    odd = pt.Context(rank=0, world_size=3)
    odd.bind(sc)
    with pytest.raises(pt.B200ptError, match="multiple of world_size"):
        odd.render_pass(1)
    odd.close()


@pytest.mark.gpu
def test_passes_on_two_ranks_sum_to_the_oracles_frame():
    """Ranks emulated one after the other on the test GPU: after every pass the reduce of the ranks' copies (here: a
    torch add into rank 0's) resolved on rank 0 is the oracle's 2-rank SPP frame at that sample count, bit for bit."""
    from owl_path_tracer_b200.distributed import accum_tensor
    W, H, world = 64, 48, 2
    sc = pt.Scene(width=W, height=H, spp=8, **CORNELL)
    ranks = []
    for r in range(world):
        c = pt.Context(rank=r, world_size=world, shard_mode=pt.SHARD_SPP)
        c.bind(sc)
        ranks.append(c)
    done = 0
    for n in (1, 3):
        infos = [c.render_pass(n) for c in ranks]
        done += n
        assert all(i["samples_done"] == done * world for i in infos)
        t0 = accum_tensor(ranks[0])
        t0 += accum_tensor(ranks[1])            # what dist.reduce(dst=0) leaves on rank 0
        ranks[0].resolve()
        o = Oracle(det=True).load_scene(pt.Scene(width=W, height=H, spp=done * world, **CORNELL))
        total = o.render(rank=0, world=world, shard_mode=pt.SHARD_SPP)["sum"] + o.render(rank=1, world=world, shard_mode=pt.SHARD_SPP)["sum"]
        want = (total[..., :3] * (np.float32(1.0) / np.float32(done * world)))[::-1]     # the frame is stored top row first
        assert (bits(ranks[0].framebuffer_f32()[..., :3]) == bits(np.ascontiguousarray(want))).all(), f"after {done} samples per rank"
        st = ranks[0].pass_stats()
        assert st["samples_done"] == done * world
        assert st["rel_variance"] == pytest.approx(rel_variance(total, done * world), rel=1e-9)
    assert infos[0]["complete"] == 1
    for c in ranks:
        c.close()
