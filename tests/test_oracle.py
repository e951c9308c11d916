"""CPU tests of the oracle itself: known-answer vectors frozen under tests/golden/ (the reference
ships none), independent re-derivations of the integer parts, analytic properties of
the BSDF restatement, and the two closest-hit finders against each other."""
import numpy as np
import pytest

import owl_path_tracer_b200 as pt
from checkers import REPO, Oracle, rel_mse

G = REPO / "tests" / "golden"


def tea_lcg(x, y, n):
    """random.hpp:46-68 re-derived in pure Python integers."""
    M = 0xFFFFFFFF
    u, v, s = x & M, y & M, 0
    for _ in range(4):
        s = (s + 0x9e3779b9) & M
        u = (u + ((((v << 4) & M) + 0xa341316c) & M ^ ((v + s) & M) ^ (((v >> 5) + 0xc8013ea4) & M))) & M
        v = (v + ((((u << 4) & M) + 0xad90777d) & M ^ ((u + s) & M) ^ (((u >> 5) + 0x7e95761e) & M))) & M
    st, out = u, []
    for _ in range(n):
        st = (16807 * st + 1013904223) & M
        out.append(np.float32(np.float32(st) * np.float32(2.0 ** -32)))
    return np.array(out, np.float32), st


@pytest.mark.parametrize("det", [False, True])
def test_rng_known_answers(det):
    o = Oracle(det=det)
    g = np.load(G / "rng.npz")
    for (x, y), bits, state in zip(g["seeds"], g["draws_bits"], g["states"]):
        d, s = o.rng(int(x), int(y), 8)
        assert (d.view(np.uint32) == bits).all()
        assert s == state
        d2, s2 = tea_lcg(int(x), int(y), 8)
        assert (d2.view(np.uint32) == bits).all() and s2 == state
    # draws lie in [0, 1]; 1.0 is reachable by rounding
    d, _ = o.rng(3, 4, 4096)
    assert d.min() >= 0.0 and d.max() <= 1.0


def test_bsdf_golden_det():
    g = np.load(G / "bsdf.npz")
    o = Oracle(det=True)
    r = o.sample_disney(g["mats"], g["material"], g["wo"], g["lobe_in"], g["rng_in"])
    assert (r["f"].view(np.uint32) == g["f_bits"]).all()
    assert (r["wi"].view(np.uint32) == g["wi_bits"]).all()
    assert (r["pdf"].view(np.uint32) == g["pdf_bits"]).all()
    assert (r["lobe"] == g["lobe_out"]).all() and (r["rng"] == g["rng_out"]).all()
    # every lobe is exercised by the fixture
    assert set(np.unique(g["lobe_out"])) >= {0, 1, 2, 3}


def test_bsdf_libm_close_to_det():
    """The glibc build is independent of the product's math; it must agree with the deterministic
    build to a few ulp wherever no branch flips."""
    g = np.load(G / "bsdf.npz")
    a = Oracle(det=False).sample_disney(g["mats"], g["material"], g["wo"], g["lobe_in"], g["rng_in"])
    same_branch = (a["lobe"] == g["lobe_out"]) & (a["rng"] == g["rng_out"])
    assert same_branch.mean() > 0.995
    f_det = g["f_bits"].view(np.float32)[same_branch]
    f_lm = a["f"][same_branch]
    ok = np.isfinite(f_det) & np.isfinite(f_lm)
    assert np.allclose(f_lm[ok], f_det[ok], rtol=2e-4, atol=1e-6)
    wi_det = g["wi_bits"].view(np.float32)[same_branch]
    assert np.allclose(a["wi"][same_branch], wi_det, rtol=0, atol=2e-4, equal_nan=True)


def test_rng_draw_counts_per_lobe():
    """Data-dependent draw counts: 1 for the lobe choice + 2 (diffuse/metal/clearcoat),
    glass 3 / 4 / 5 after the lobe draw (Q11)."""
    g = np.load(G / "bsdf.npz")
    A, Cc, M = 16807, 1013904223, 2 ** 32

    def advance(s, k):
        for _ in range(k):
            s = (A * s + Cc) % M
        return s
    for i in range(0, g["material"].size, 5):
        s0, s1, lobe = int(g["rng_in"][i]), int(g["rng_out"][i]), int(g["lobe_out"][i])
        counts = [k for k in range(0, 8) if advance(s0, k) == s1]
        assert counts, "rng state is not on the stream"
        k = counts[0]
        if lobe in (0, 1, 2):
            assert k == 3
        elif lobe == 3:
            assert k in (4, 5, 6)  # transmitted | TIR (no Fresnel draw) | reflected


def test_diffuse_white_furnace():
    """Lambert limit: roughness 0, white base colour -> f*cos/pdf == 1 - (schlick terms); mean weight over
    many samples stays <= 1 and close to the analytic value of the Disney diffuse lobe."""
    o = Oracle(det=False)
    m = np.zeros(1, pt.MATERIAL_DTYPE)
    for k in ("base_color_r", "base_color_g", "base_color_b"):
        m[0][k] = 1.0
    m[0]["roughness"] = 0.0
    m[0]["ior"] = 1.5
    n = 20000
    rs = np.random.RandomState(1)
    wo = np.tile(np.array([[0, 0, 1]], np.float32), (n, 1))
    r = o.sample_disney(m, np.zeros(n, np.int32), wo, np.full(n, -1, np.int32), rs.randint(0, 2 ** 32, n, dtype=np.uint64).astype(np.uint32))
    assert (r["lobe"] == 0).all()
    w = r["f"][:, 0] * np.abs(r["wi"][:, 2]) / r["pdf"]
    ok = r["pdf"] >= 1e-5
    # normal incidence: Fo = 0, weight = 1 - 0.5*Fi
    fi = np.clip(1 - r["wi"][ok, 2], 0, 1) ** 5
    assert np.allclose(w[ok], 1 - 0.5 * fi, atol=2e-5)
    assert 0.9 < w[ok].mean() <= 1.0
    # cosine-hemisphere sampling: E[cos] = 2/3
    assert abs(r["wi"][ok, 2].mean() - 2 / 3) < 0.01


def test_camera_matches_host_loader():
    """camera.cpp:3-22 restated in the oracle == the host loader's camera (both glibc)."""
    o = Oracle(det=False)
    sc = pt.Scene(scene="cornell-box", width=512, height=512, spp=1, depth=1)
    cam = o.camera([3, 1, 0], [0, 1, 0], [0, 1, 0], 50.0, 512, 512)
    for a, b in zip(cam, sc.camera):
        assert (a.view(np.uint32) == b.view(np.uint32)).all()
    g = np.load(G / "camera.npz")
    assert np.allclose(cam, g["cam_bits"].view(np.float32), rtol=1e-6, atol=1e-7)


def test_quantise_make_rgba():
    """owl_device.h:79-91: int(f*256) clamped, NaN -> 0 (device semantics)."""
    o = Oracle()
    q = o.L.oracle_quantise
    assert q(0.0, 0.5, 1.0) == (0 | (128 << 8) | (255 << 16) | (255 << 24))
    assert q(-1.0, 0.999, 100.0) == (0 | (255 << 8) | (255 << 16) | (255 << 24))
    assert q(float("nan"), float("inf"), -float("inf")) == (0 | (255 << 8) | (0 << 16) | (255 << 24))
    assert q(1 / 256, 1.5 / 256, 254.99 / 256) == (1 | (1 << 8) | (254 << 16) | (255 << 24))


@pytest.fixture(scope="module")
def cornell64():
    sc = pt.Scene(scene="cornell-box", width=64, height=64, spp=16, depth=8, env="color", env_intensity=0.0)
    return sc, Oracle(det=True).load_scene(sc)


def test_cornell_golden(cornell64):
    sc, o = cornell64
    g = np.load(G / "cornell_64.npz")
    h = o.primary_hits()
    assert (h["mesh"] == g["mesh"]).all() and (h["prim"] == g["prim"]).all()
    assert (h["t"].view(np.uint32) == g["t_bits"]).all()
    img = o.render()
    assert (img["sum"].view(np.uint32)[..., :3] == g["sum_bits"][..., :3]).all()   # .w: second moment, added after the fixture was frozen
    assert (img["rgba"] == g["rgba"]).all()
    assert img["rays"] == int(g["rays"])


def test_bvh_finder_equals_brute_force(cornell64):
    sc, o = cornell64
    a, b = o.primary_hits(brute=False), o.primary_hits(brute=True)
    for k in ("mesh", "prim"):
        assert (a[k] == b[k]).all()
    for k in ("t", "u", "v"):
        assert (a[k].view(np.uint32) == b[k].view(np.uint32)).all()
    # incoherent rays from inside the box
    rs = np.random.RandomState(3)
    n = 600
    org = np.tile(np.array([[0.3, 1.0, 0.1]], np.float32), (n, 1)) + rs.uniform(-.2, .2, (n, 3)).astype(np.float32)
    d = rs.normal(size=(n, 3)).astype(np.float32)
    d /= np.linalg.norm(d, axis=1, keepdims=True)
    x, y = o.trace(org, d, brute=False), o.trace(org, d, brute=True)
    assert (x["mesh"] == y["mesh"]).all() and (x["prim"] == y["prim"]).all()
    assert (x["t"].view(np.uint32) == y["t"].view(np.uint32)).all()


@pytest.mark.parametrize("name", ["cube", "sphere"])
def test_small_scene_golden(name):
    kw = dict(cube=dict(env="auto", env_intensity=1.0), sphere=dict(env="map", env_intensity=1.0))[name]
    sc = pt.Scene(scene=name, width=48, height=48, spp=8, depth=8, **kw)
    img = Oracle(det=True).load_scene(sc).render()
    g = np.load(G / f"{name}_48.npz")
    assert (img["sum"].view(np.uint32)[..., :3] == g["sum_bits"][..., :3]).all()   # .w: second moment, not in the frozen fixture
    assert (img["rgba"] == g["rgba"]).all()
    assert img["rays"] == int(g["rays"])


def test_libm_and_det_images_agree_statistically():
    """Independent-math oracle vs deterministic-math oracle: same estimator, paths diverge after a
    few bounces, images agree within Monte-Carlo noise (relMSE of two 256-spp renders)."""
    sc = pt.Scene(scene="cornell-box", width=32, height=32, spp=256, depth=8, env="color", env_intensity=0.0)
    a = Oracle(det=False).load_scene(sc).render()["mean"]
    b = Oracle(det=True).load_scene(sc).render()["mean"]
    assert rel_mse(a, b) < 5e-3
    assert abs(a[..., :3].mean() - b[..., :3].mean()) < 0.01 * b[..., :3].mean()


def test_sharding_keys_of_the_oracle():
    """rank/world keying: ROWS shards tile the 1-rank image exactly; SPP shards use
    per-rank streams, rank 0's stream being the reference stream."""
    sc = pt.Scene(scene="cornell-box", width=24, height=24, spp=6, depth=4, env="color", env_intensity=0.0)
    o = Oracle(det=True).load_scene(sc)
    full = o.render()["sum"]
    rows = sum(o.render(rank=r, world=3, shard_mode=1)["sum"] for r in range(3))
    assert (rows.view(np.uint32) == full.view(np.uint32)).all()
    s0 = o.render(rank=0, world=2, shard_mode=0)["sum"]
    sc3 = pt.Scene(scene="cornell-box", width=24, height=24, spp=3, depth=4, env="color", env_intensity=0.0)
    first_half = Oracle(det=True).load_scene(sc3).render()["sum"]
    assert (s0.view(np.uint32) == first_half.view(np.uint32)).all()
    s1 = o.render(rank=1, world=2, shard_mode=0)["sum"]
    assert not (s1 == s0).all()
