"""CPU parity: the product's per-lane / per-element functions (compiled for the host by
tests/emu) against the scalar oracle.  Bit-exact: hits (mesh, prim, t, u, v) and float sums.
These run without a GPU; tests/test_gpu_parity.py repeats them on the device."""
import numpy as np
import pytest

import owl_path_tracer_b200 as pt
from checkers import REPO, Emu, Oracle, unpack_pairs

G = REPO / "tests" / "golden"
SCENES = {
    "cornell-box": dict(env="color", env_intensity=0.0),
    "cube": dict(env="auto", env_intensity=1.0),
    "sphere": dict(env="map", env_intensity=1.0),
    "mitsuba": dict(env="auto", env_intensity=1.0),
    "car": dict(env="color", env_intensity=0.0, detail=1),
    "dragon": dict(env="color", env_intensity=0.0, detail=60000),
}


def bits(a):
    return np.ascontiguousarray(a).view(np.uint32)


def same_hits(a, b):
    return ((a["mesh"] == b["mesh"]) & (a["prim"] == b["prim"]) & (bits(a["t"]) == bits(b["t"])) &
            (bits(a["u"]) == bits(b["u"])) & (bits(a["v"]) == bits(b["v"])))


@pytest.mark.parametrize("name", list(SCENES))
def test_scene_bit_parity(name):
    sc = pt.Scene(scene=name, width=48, height=36, spp=4, depth=8, **SCENES[name])
    o, e = Oracle(det=True).load_scene(sc), Emu().load_scene(sc)
    assert same_hits(o.primary_hits(), e.primary_hits()).all()
    ro, re = o.render(), e.render()
    assert ro["rays"] == re["rays"]
    assert (bits(ro["sum"]) == bits(re["sum"])).all()
    assert (ro["rgba"] == re["rgba"]).all()
    assert (bits(ro["mean"]) == bits(re["mean"])).all()


def test_bsdf_matches_golden():
    g = np.load(G / "bsdf.npz")
    r = Emu().sample_disney(g["mats"], g["material"], g["wo"], g["lobe_in"], g["rng_in"])
    assert (bits(r["f"]) == g["f_bits"]).all() and (bits(r["wi"]) == g["wi_bits"]).all()
    assert (bits(r["pdf"]) == g["pdf_bits"]).all()
    assert (r["lobe"] == g["lobe_out"]).all() and (r["rng"] == g["rng_out"]).all()


def test_material_sweep_all_lobes():
    """mitsuba material test (BASELINE config 3): sweep the attributes settings.json's `test`
    block is meant for; every frame bit-equal to the oracle."""
    sc = pt.Scene(scene="mitsuba", width=40, height=24, spp=3, depth=8, env="auto", env_intensity=1.0, detail=0)
    o, e = Oracle(det=True).load_scene(sc), Emu().load_scene(sc)
    for attr, val in (("metallic", 1.0), ("roughness", 0.1), ("clearcoat", 1.0), ("specular_transmission", 1.0),
                      ("sheen", 1.0), ("anisotropic", 0.8)):
        sc.set_material("outside", attr, val)
        o.set_materials(sc.materials())
        e.set_materials(sc.materials())
        ro, re = o.render(), e.render()
        assert (bits(ro["sum"]) == bits(re["sum"])).all(), attr
        assert ro["rays"] == re["rays"]


def random_rays(n, seed, centre, spread):
    rs = np.random.RandomState(seed)
    o = (np.asarray(centre, np.float32) + rs.uniform(-spread, spread, (n, 3))).astype(np.float32)
    d = rs.normal(size=(n, 3)).astype(np.float32)
    d /= np.linalg.norm(d, axis=1, keepdims=True)
    return o, d


def test_incoherent_rays_vs_brute_force():
    sc = pt.Scene(scene="cornell-box", width=8, height=8, spp=1, depth=1)
    o, e = Oracle(det=True).load_scene(sc), Emu().load_scene(sc)
    org, d = random_rays(1500, 5, [0.2, 1.0, 0.0], 0.6)
    # axis-parallel and zero-component directions stress the slab test
    d[:50] = np.eye(3, dtype=np.float32)[np.arange(50) % 3] * np.where(np.arange(50) % 2, 1, -1)[:, None]
    d[50:100, 1] = 0
    d[50:100] /= np.linalg.norm(d[50:100], axis=1, keepdims=True)
    assert same_hits(o.trace(org, d, brute=True), e.trace(org, d)).all()
    assert e.last_counters[0] > 0 and e.last_counters[1] > 0


def make_soup(n, seed, degenerate=False):
    rs = np.random.RandomState(seed)
    c = rs.uniform(-1, 1, (n, 1, 3))
    v = (c + rs.normal(scale=0.15, size=(n, 3, 3))).astype(np.float32)
    if degenerate:
        v[::5, 2] = v[::5, 1]          # zero-area triangles
        v[1::7] = v[0:1]               # many copies of the same triangle (equal Morton keys, t ties)
    verts = v.reshape(-1, 3)
    normals = np.tile(np.array([[0, 1, 0]], np.float32), (verts.shape[0], 1))
    idx = np.arange(3 * n, dtype=np.int32).reshape(n, 3)
    return verts, normals, idx


@pytest.mark.parametrize("n,degenerate", [(1, False), (2, False), (3, False), (4, False), (9, False), (257, True), (5000, True)])
def test_triangle_soups_and_edge_counts(n, degenerate):
    """Single triangle, tiny counts (leaf-only roots), duplicates and zero-area triangles: the
    builder must keep every triangle exactly once and the walk must agree with brute force,
    including the tie rule (equal t -> smaller id)."""
    verts, normals, idx = make_soup(n, 100 + n, degenerate)
    mats = np.zeros(1, pt.MATERIAL_DTYPE)
    o, e = Oracle(det=True), Emu()
    for c in (o, e):
        c.add_mesh(verts, normals, idx, 0)
        c.set_materials(mats)
        c.build()
    nodes, tris = e.export()
    ids = np.sort(unpack_pairs(tris)[0])
    assert (ids == np.arange(n)).all()
    org, d = random_rays(800, n, [0, 0, 0], 1.5)
    assert same_hits(o.trace(org, d, brute=True), e.trace(org, d)).all()


def surface_rays(sc, n, seed):
    """Bounce-like rays: origins ON triangles (so they sit on leaf and node box faces), directions
    random, grazing the surface, or with exactly zero components."""
    rs = np.random.RandomState(seed)
    T = np.concatenate([en["verts"][en["idx"]] for en in sc.entities()]).astype(np.float32)
    k = rs.randint(0, T.shape[0], n)
    a, b = rs.rand(n, 1).astype(np.float32), rs.rand(n, 1).astype(np.float32)
    f = (a + b > 1).ravel()
    a[f], b[f] = 1 - a[f], 1 - b[f]
    P = (T[k, 0] * (1 - a - b) + T[k, 1] * a + T[k, 2] * b).astype(np.float32)
    P[: n // 8] = T[k[: n // 8], 0]                       # exactly on a vertex
    D = rs.normal(size=(n, 3)).astype(np.float32)
    e1, e2 = T[k, 1] - T[k, 0], T[k, 2] - T[k, 0]
    g = slice(n // 4, n // 2)                              # grazing: in-plane direction + tiny lift
    nrm = np.cross(e1[g], e2[g])
    D[g] = e1[g] * rs.normal(size=(n // 2 - n // 4, 1)) + e2[g] * rs.normal(size=(n // 2 - n // 4, 1)) + 1e-4 * nrm
    z = slice(n // 2, 3 * n // 4)                          # one or two exactly-zero components
    D[z, rs.randint(0, 3, 3 * n // 4 - n // 2)] = 0
    D[n // 2: n // 2 + n // 16, 1] = 0
    D[~np.isfinite(D).all(axis=1) | (np.abs(D).sum(axis=1) == 0)] = [0, 1, 0]
    D = (D / np.linalg.norm(D, axis=1, keepdims=True)).astype(np.float32)
    return np.ascontiguousarray(P), np.ascontiguousarray(D)


@pytest.mark.parametrize("name", ["cornell-box", "cube", "sphere"])
def test_surface_rays_vs_brute_force(name):
    """Rays that start on the geometry stress the conservativeness of the quantised boxes (a hit at
    t just above t_min lies on a box face): the walk must never cull what brute force finds."""
    sc = pt.Scene(scene=name, width=8, height=8, spp=1, depth=1)
    o, e = Oracle(det=True).load_scene(sc), Emu().load_scene(sc)
    P, D = surface_rays(sc, 40000 if name != "cornell-box" else 12000, 11)
    hb, he = o.trace(P, D, brute=True), e.trace(P, D)
    bad = np.argwhere(~same_hits(hb, he)).ravel()
    # The only tolerated disagreement (the "documented tie" of BASELINE.md section 3): a ray lying
    # almost IN the plane of a triangle makes Moller-Trumbore ill-conditioned and brute force can
    # report a hit whose point o + t d is outside that triangle's bounding box.  No bounding volume
    # hierarchy can return such a phantom hit (the oracle's own BVH drops it as well).
    ents = sc.entities()
    max_abs = max(np.abs(en["verts"]).max() for en in ents)
    for i in bad:
        m, k = (hb[i]["mesh"], hb[i]["prim"]) if hb[i]["mesh"] >= 0 else (he[i]["mesh"], he[i]["prim"])
        t = hb[i]["t"] if hb[i]["mesh"] >= 0 else he[i]["t"]
        tri = ents[m]["verts"][ents[m]["idx"][k]].astype(np.float64)
        hp = P[i].astype(np.float64) + float(t) * D[i].astype(np.float64)
        pad = max_abs * 2.0 ** -18
        outside = ((hp < tri.min(0) - pad) | (hp > tri.max(0) + pad)).any()
        assert outside, f"ray {i}: walk disagrees with brute force on a well-conditioned hit"
    assert len(bad) <= len(P) // 2000


def test_fallback_radix_tree_builder():
    """The builder falls back to the Morton radix tree when the clustered tree is deeper than the
    traversal stack; force that path and check it against brute force too."""
    verts, normals, idx = make_soup(3000, 77, True)
    mats = np.zeros(1, pt.MATERIAL_DTYPE)
    o, e = Oracle(det=True), Emu()
    e.L.emu_force_lbvh(e.h, 1)
    for c in (o, e):
        c.add_mesh(verts, normals, idx, 0)
        c.set_materials(mats)
        c.build()
    org, d = random_rays(800, 5, [0, 0, 0], 1.5)
    assert same_hits(o.trace(org, d, brute=True), e.trace(org, d)).all()


def decode_nodes(nodes):
    """80-byte nodes -> origin, scales, imask, child_base, tri_base, meta, lo[3][8], hi[3][8]"""
    p = nodes[:, 0:12].view(np.float32)
    e = nodes[:, 12:15].astype(np.uint32)
    scale = (e << 23).view(np.float32)
    imask = nodes[:, 15]
    base = nodes[:, 16:24].view(np.uint32)
    meta = nodes[:, 24:32]
    q = nodes[:, 32:80].reshape(-1, 6, 8).astype(np.float32)
    lo = p[:, :, None] + q[:, 0:3] * scale[:, :, None]
    hi = p[:, :, None] + q[:, 3:6] * scale[:, :, None]
    return p, scale, imask, base[:, 0], base[:, 1], meta, lo, hi


def test_tree_invariants():
    """Every leaf's quantised box contains its triangles; every inner child's box contains the
    child's node origin box; inner children are numbered by slot order; each triangle once."""
    sc = pt.Scene(scene="cornell-box", width=8, height=8, spp=1, depth=1)
    e = Emu().load_scene(sc)
    nodes, tris = e.export()
    ids, v0, e1, e2, blk = unpack_pairs(tris)
    assert (np.sort(ids) == np.arange(sc.n_triangles)).all()
    p, scale, imask, cbase, tbase, meta, lo, hi = decode_nodes(nodes)
    corners = np.stack([v0, v0 + e1, v0 + e2], 1)
    seen_nodes = np.zeros(e.n_nodes, bool)
    seen_nodes[0] = True
    for n in range(e.n_nodes):
        rank = 0
        for s in range(8):
            m = int(meta[n, s])
            if m == 0:
                continue
            if (m & 0x18) == 0x18 and (m >> 5) == 1:
                assert (m & 0x1F) == 24 + s and (imask[n] >> s) & 1
                child = int(cbase[n]) + rank
                rank += 1
                assert not seen_nodes[child]
                seen_nodes[child] = True
                # the child's grid origin sits 1/8 step below the child's box (collapse_node)
                assert (lo[n, :, s] <= p[child] + 0.125 * scale[child] + 1e-6).all()
            else:
                cnt = {1: 1, 3: 2, 7: 3}[m >> 5]                      # pair blocks of the leaf
                first = int(tbase[n]) + (m & 0x1F)
                sel = (blk >= first) & (blk < first + cnt)
                assert cnt * 2 - 1 <= sel.sum() <= cnt * 2            # only the last block may be half empty
                c = corners[sel].reshape(-1, 3)
                assert (c >= lo[n, :, s] - 1e-7).all() and (c <= hi[n, :, s] + 1e-7).all()
        assert rank == bin(int(imask[n])).count("1")
    assert seen_nodes.all()
    assert e.levels <= 32


def test_reinsertion_passes_lower_the_tree_area():
    """pt_build.h "reinsertion optimisation": the passes between clustering and collapse move nodes, the
    binary tree's surface area drops, and the finished tree still holds every triangle once (the closest
    hits of test_surface_rays_vs_brute_force and the invariants above run on the tree they leave)."""
    sc = pt.Scene(scene="cornell-box", width=8, height=8, spp=1, depth=1)
    e = Emu().load_scene(sc)
    moves, before, after = e.reinsert_stats()
    assert moves > 100
    assert after < before * 0.995
    ids = unpack_pairs(e.export()[1])[0]
    assert (np.sort(ids) == np.arange(sc.n_triangles)).all()


def test_sharded_renders_match_oracle_keys():
    sc = pt.Scene(scene="cornell-box", width=20, height=18, spp=5, depth=5, env="color", env_intensity=0.0)
    o, e = Oracle(det=True).load_scene(sc), Emu().load_scene(sc)
    full = e.render()["sum"]
    for world in (2, 3):
        rows = np.zeros_like(full)
        for r in range(world):
            a, b = e.render(rank=r, world=world, shard_mode=1), o.render(rank=r, world=world, shard_mode=1)
            assert (bits(a["sum"]) == bits(b["sum"])).all()
            rows += a["sum"]
        assert (bits(rows) == bits(full)).all()   # ROWS: the reduce of the shards IS the 1-rank frame
        for r in range(world):
            a, b = e.render(rank=r, world=world, shard_mode=0), o.render(rank=r, world=world, shard_mode=0)
            assert (bits(a["sum"]) == bits(b["sum"])).all()
            assert a["rays"] == b["rays"]
