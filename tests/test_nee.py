"""Next-event estimation mode (B200PT_FLAG_NEE): light sampling of the emissive
triangles + shadow any-hit ray + power-heuristic MIS.

Lights are the emissive triangles and, when an environment map is in use, the map (texels by luminance).
The reference holds the ingredients but never calls them (lights.hpp sample_light / power_heuristic,
visibiliy_test device.cu:97-110), so there is no reference image of this mode: the oracle carries the
same opt-in estimator (oracle_set_nee) and parity is product == oracle, bit for bit, as for the
reference's estimator.  Against the reference's estimator the mode is checked statistically: same
expectation at the depth where both see the same light paths, lower variance at equal spp."""
import numpy as np
import pytest

import owl_path_tracer_b200 as pt
from checkers import Emu, Oracle, rel_mse

SCENES = {
    "cornell-box": dict(env="color", env_intensity=0.0),
    "sphere": dict(env="map", env_intensity=1.0),          # no emissive triangle: the environment map is the light
    "mitsuba": dict(env="auto", env_intensity=1.0),        # analytic sky, no emissive triangle: nothing to sample
    "car": dict(env="color", env_intensity=0.0, detail=1),
}


def bits(a):
    return np.ascontiguousarray(a).view(np.uint32)


@pytest.mark.parametrize("name", list(SCENES))
def test_nee_emulation_bit_equals_oracle(name):
    sc = pt.Scene(scene=name, width=48, height=36, spp=4, depth=8, **SCENES[name])
    ro = Oracle(det=True).load_scene(sc).set_nee().render()
    re = Emu().load_scene(sc).set_nee().render()
    assert ro["rays"] == re["rays"] and ro["shadow_rays"] == re["shadow_rays"]
    assert (bits(ro["sum"]) == bits(re["sum"])).all()
    assert (ro["rgba"] == re["rgba"]).all()
    if name == "mitsuba":
        assert ro["shadow_rays"] == 0
    else:
        assert ro["shadow_rays"] > 0


def test_nee_switch_off_restores_reference_estimator():
    sc = pt.Scene(scene="cornell-box", width=32, height=32, spp=2, depth=8, env="color", env_intensity=0.0)
    a = Oracle(det=True).load_scene(sc).render()
    o = Oracle(det=True).load_scene(sc)
    o.set_nee(True).render()
    b = o.set_nee(False).render()
    assert a["rays"] == b["rays"] and b["shadow_rays"] == 0
    assert (bits(a["sum"]) == bits(b["sum"])).all()


def test_nee_same_expectation_lower_variance():
    # depth 8, closed box: the two estimators integrate the same light paths up to the last bounce
    kw = dict(scene="cornell-box", width=40, height=40, depth=8, env="color", env_intensity=0.0)
    truth = Oracle(det=False).load_scene(pt.Scene(spp=1024, **kw)).set_nee().render()["mean"][..., :3]
    sc = pt.Scene(spp=16, **kw)
    ref = Oracle(det=False).load_scene(sc).render()["mean"][..., :3]
    nee = Oracle(det=False).load_scene(sc).set_nee().render()["mean"][..., :3]
    long_ref = Oracle(det=False).load_scene(pt.Scene(spp=512, **kw)).render()["mean"][..., :3]
    # expectation: frame means of the two estimators agree within 3 %
    assert abs(long_ref.mean() - truth.mean()) <= 0.03 * truth.mean()
    # variance: at 16 spp NEE is several times closer to the converged frame
    e_ref, e_nee = rel_mse(ref, truth), rel_mse(nee, truth)
    assert e_nee < 0.5 * e_ref, (e_nee, e_ref)


def test_nee_map_and_triangles_together():
    """A scene with an emissive quad AND an environment map in use: one extra draw picks the strategy."""
    sc = pt.Scene(scene="cornell-box", width=40, height=30, spp=4, depth=6, env="map", env_intensity=1.0)
    assert sc.environment["rgba"].size > 0 and sc.environment["use"]
    ro = Oracle(det=True).load_scene(sc).set_nee().render()
    re = Emu().load_scene(sc).set_nee().render()
    assert ro["rays"] == re["rays"] and ro["shadow_rays"] == re["shadow_rays"] and ro["shadow_rays"] > 0
    assert (bits(ro["sum"]) == bits(re["sum"])).all()


def test_nee_environment_map_same_expectation():
    kw = dict(scene="sphere", width=48, height=36, depth=8, env="map", env_intensity=1.0)
    ref = Oracle(det=False).load_scene(pt.Scene(spp=256, **kw)).render()["mean"][..., :3]
    nee = Oracle(det=False).load_scene(pt.Scene(spp=256, **kw)).set_nee().render()["mean"][..., :3]
    assert abs(ref.mean() - nee.mean()) <= 0.01 * ref.mean()
    assert rel_mse(nee, ref) < 2e-3          # two independent 256-spp estimates of the same image


def hdr_sky(w=64, h=32, sun=400.0):
    """Small float environment: dim gradient sky and a 2x2-texel sun 400x brighter (what the reference's LDR
    route cannot represent) in the upper hemisphere (v > 0.5 is up: v = 0.5 + asin(y)/pi)."""
    v = (np.arange(h, dtype=np.float32)[:, None] + 0.5) / h
    img = np.empty((h, w, 3), np.float32)
    img[...] = (0.05 + 0.25 * v)[..., None] * np.array([0.6, 0.8, 1.0], np.float32)
    img[h * 3 // 4: h * 3 // 4 + 2, w // 3: w // 3 + 2] = sun
    return img


def test_float_environment_emulation_bit_equals_oracle():
    sc = pt.Scene(scene="sphere", width=48, height=36, spp=4, depth=8, env="map", env_intensity=1.0)
    sky = hdr_sky()
    for nee in (False, True):
        o, e = Oracle(det=True).load_scene(sc), Emu().load_scene(sc)
        for c in (o, e):
            c.set_environment_f32(sky, use=True, intensity=1.0)
            c.set_nee(nee)
        ro, re = o.render(), e.render()
        assert ro["rays"] == re["rays"] and ro["shadow_rays"] == re["shadow_rays"]
        assert (bits(ro["sum"]) == bits(re["sum"])).all()
        if nee:  # radiance above what an LDR map can carry reaches the frame
            assert ro["mean"][..., :3].max() > 1.0
        assert (ro["shadow_rays"] > 0) == nee


def test_host_loader_float_environment_route():
    """Scene(env_float=True): environment.hdr kept as linear radiance next to the reference's 8-bit conversion
    (same pixels, same row order), bound through b200pt_set_environment_f32 by the C++ host layer."""
    sc = pt.Scene(scene="sphere", width=48, height=36, spp=4, depth=8, env="map", env_intensity=1.0, env_float=True)
    env = sc.environment
    f, ldr = env["rgb_f32"], env["rgba"]
    assert f is not None and f.shape[:2] == ldr.shape[:2] and f.max() > 4.0        # the sun is far above 1
    want = np.clip(np.power(np.maximum(f.astype(np.float64), 0), 1 / 2.2) * 255 + 0.5, 0, 255).astype(np.uint8)   # stb_image.h:1864-1889
    assert (np.abs(want.astype(int) - ldr[..., :3].astype(int)) <= 1).all()
    ro = Oracle(det=True).load_scene(sc).set_nee().render()
    re = Emu().load_scene(sc).set_nee().render()
    assert ro["rays"] == re["rays"] and ro["shadow_rays"] == re["shadow_rays"] and ro["shadow_rays"] > 0
    assert (bits(ro["sum"]) == bits(re["sum"])).all()
    assert pt.Scene(scene="sphere", width=8, height=8, spp=1, depth=1, env="map").environment["rgb_f32"] is None


def test_float_environment_importance_sampling_pays():
    """With a small bright sun, light sampling of the map (texels by luminance) + MIS must beat BSDF sampling
    alone at equal spp, and both must agree in expectation."""
    sc = pt.Scene(scene="sphere", width=40, height=30, spp=32, depth=4, env="map", env_intensity=1.0)
    sky = hdr_sky()

    def run(nee, spp):
        o = Oracle(det=False).load_scene(pt.Scene(scene="sphere", width=40, height=30, spp=spp, depth=4, env="map", env_intensity=1.0))
        o.set_environment_f32(sky, use=True, intensity=1.0)
        return o.set_nee(nee).render()["mean"][..., :3]
    truth = run(True, 2048)
    ref, nee = run(False, 32), run(True, 32)
    long_ref = run(False, 2048)
    assert abs(long_ref.mean() - truth.mean()) <= 0.05 * truth.mean()
    e_ref, e_nee = rel_mse(ref, truth), rel_mse(nee, truth)
    assert e_nee < 0.25 * e_ref, (e_nee, e_ref)


@pytest.mark.gpu
def test_host_loader_float_environment_gpu():
    sc = pt.Scene(scene="sphere", width=160, height=120, spp=4, depth=8, env="map", env_intensity=1.0, env_float=True)
    for flags in (0, pt.FLAG_NEE):
        ro = Oracle(det=True).load_scene(sc).set_nee(flags != 0).render()
        ctx = pt.Context(flags=flags)
        ctx.bind(sc)                      # C++ host layer -> b200pt_set_environment_f32
        ctx.render()
        assert ctx.stats()["rays"] == ro["rays"] and ctx.stats()["shadow_rays"] == ro["shadow_rays"]
        assert (bits(ctx.framebuffer_f32()[..., :3]) == bits(ro["mean"][..., :3])).all()
        ctx.close()


@pytest.mark.gpu
def test_float_environment_gpu_bit_equals_oracle():
    sc = pt.Scene(scene="sphere", width=160, height=120, spp=4, depth=8, env="map", env_intensity=1.0)
    sky = hdr_sky(128, 64)
    for flags in (0, pt.FLAG_NEE):
        o = Oracle(det=True).load_scene(sc)
        o.set_environment_f32(sky, use=True, intensity=1.0)
        ro = o.set_nee(flags != 0).render()
        ctx = pt.Context(flags=flags)
        ctx.bind(sc)
        ctx.set_environment_f32(sky, use=True, intensity=1.0)
        ctx.render()
        st = ctx.stats()
        assert st["rays"] == ro["rays"] and st["shadow_rays"] == ro["shadow_rays"]
        assert (bits(ctx.framebuffer_f32()[..., :3]) == bits(ro["mean"][..., :3])).all()
        # back to the RGBA8 map: the float map is gone
        env = sc.environment
        ctx.set_environment(env["rgba"], env["use"], env["auto"], env["color"], env["intensity"])
        ctx.render()
        o2 = Oracle(det=True).load_scene(sc).set_nee(flags != 0).render()
        assert (bits(ctx.framebuffer_f32()[..., :3]) == bits(o2["mean"][..., :3])).all()
        ctx.close()


@pytest.mark.gpu
@pytest.mark.parametrize("name", ["cornell-box", "sphere", "mitsuba", "car", "dragon"])
def test_nee_gpu_bit_equals_oracle(name):
    kw = dict(SCENES.get(name, dict(env="color", env_intensity=0.0)))
    kw.pop("detail", None)
    sc = pt.Scene(scene=name, width=160, height=90, spp=4, depth=8, **kw)
    ro = Oracle(det=True).load_scene(sc).set_nee().render()
    ctx = pt.Context(flags=pt.FLAG_NEE)
    ctx.bind(sc)
    ctx.render()
    st = ctx.stats()
    assert st["rays"] == ro["rays"] and st["shadow_rays"] == ro["shadow_rays"]
    assert (bits(ctx.framebuffer_f32()[..., :3]) == bits(ro["mean"][..., :3])).all()
    assert (ctx.framebuffer_rgba8() == ro["rgba"]).all()
    # a second frame on the same context starts from clean state
    ctx.render()
    assert (bits(ctx.framebuffer_f32()[..., :3]) == bits(ro["mean"][..., :3])).all()
    ctx.close()


@pytest.mark.gpu
def test_nee_gpu_map_and_triangles_together():
    sc = pt.Scene(scene="cornell-box", width=160, height=120, spp=4, depth=8, env="map", env_intensity=1.0)
    ro = Oracle(det=True).load_scene(sc).set_nee().render()
    ctx = pt.Context(flags=pt.FLAG_NEE)
    ctx.bind(sc)
    ctx.render()
    st = ctx.stats()
    assert st["rays"] == ro["rays"] and st["shadow_rays"] == ro["shadow_rays"] and st["shadow_rays"] > 0
    assert (bits(ctx.framebuffer_f32()[..., :3]) == bits(ro["mean"][..., :3])).all()
    ctx.close()


@pytest.mark.gpu
def test_nee_gpu_cornell_frame_and_sharding():
    sc = pt.Scene(scene="cornell-box", width=128, height=128, spp=8, depth=8, env="color", env_intensity=0.0)
    ro = Oracle(det=True).load_scene(sc).set_nee().render()
    # rows sharding, both ranks on one GPU: sums add up to the single-GPU frame bit for bit
    from owl_path_tracer_b200.distributed import accum_tensor
    total = None
    shadow = 0
    for r in range(2):
        ctx = pt.Context(rank=r, world_size=2, shard_mode=pt.SHARD_ROWS, flags=pt.FLAG_NEE)
        ctx.bind(sc)
        ctx.render()
        t = accum_tensor(ctx).clone()
        total = t if total is None else total + t
        shadow += ctx.stats()["shadow_rays"]
        ctx.close()
    assert shadow == ro["shadow_rays"]
    assert (bits(total.cpu().numpy().reshape(128, 128, 4)[..., :3]) == bits(ro["sum"][..., :3])).all()
