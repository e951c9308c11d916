"""ctypes wrappers of the CHECKERS used by the tests (never by the product):

* ``Oracle``  oracle/liboracle.so (glibc libm) or oracle/liboracle_det.so (deterministic math):
  scalar restatement of the reference integrator.
* ``Emu``     tests/emu/libemu.so: the product's own per-lane / per-element functions compiled
  for the host, run in plain loops (no GPU needed).
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess
from pathlib import Path

import numpy as np

REPO = Path(__file__).resolve().parent.parent
HIT_DTYPE = np.dtype([("mesh", np.int32), ("prim", np.int32), ("t", np.float32), ("u", np.float32), ("v", np.float32)])
NCPU = max(1, min(os.cpu_count() or 1, 16))


def _p(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


def ensure_built():
    need = [REPO / "oracle/liboracle.so", REPO / "oracle/liboracle_det.so", REPO / "tests/emu/libemu.so"]
    if not all(p.exists() for p in need):
        subprocess.run(["make", "-s", "oracle", "emu"], cwd=REPO, check=True)


class _SceneChecker:
    prefix = ""

    def _f(self, name):
        return getattr(self.L, self.prefix + name)

    def load_scene(self, scene):
        """Feed a owl_path_tracer_b200.Scene (host loader output) into the checker."""
        for e in scene.entities():
            self.add_mesh(e["verts"], e["normals"], e["idx"], e["material"], e["uvs"], e["tex"])
        self.set_materials(scene.materials())
        env = scene.environment
        if env.get("rgb_f32") is not None:
            self.set_environment_f32(env["rgb_f32"], env["use"], env["auto"], env["color"], env["intensity"])
        else:
            self.set_environment(env["rgba"], env["use"], env["auto"], env["color"], env["intensity"])
        self.set_camera(*scene.camera)
        self.set_render(scene.width, scene.height, scene.spp, scene.depth)
        self.build()
        return self

    def add_mesh(self, verts, normals, idx, material, uvs=None, tex=None):
        verts = np.ascontiguousarray(verts, np.float32)
        normals = np.ascontiguousarray(normals, np.float32)
        idx = np.ascontiguousarray(idx, np.int32)
        uvs = None if uvs is None else np.ascontiguousarray(uvs, np.float32)
        tex = None if tex is None else np.ascontiguousarray(tex, np.uint8)
        th, tw = (tex.shape[0], tex.shape[1]) if tex is not None else (0, 0)
        r = self._f("add_mesh")(self.h, _p(verts), _p(normals), _p(uvs), verts.shape[0], _p(idx), idx.shape[0],
                                int(material), _p(tex), tw, th)
        assert r >= 0
        return r

    def set_materials(self, mats):
        m = np.ascontiguousarray(mats).view(np.float32).reshape(-1, 17)
        self.mats = np.ascontiguousarray(m)
        self._f("set_materials")(self.h, _p(self.mats), self.mats.shape[0])

    def set_environment(self, rgba=None, use=False, auto=False, color=(0, 0, 0), intensity=0.0):
        rgba = None if rgba is None else np.ascontiguousarray(rgba, np.uint8)
        h, w = (rgba.shape[0], rgba.shape[1]) if rgba is not None else (0, 0)
        col = np.asarray(color, np.float32)
        self._f("set_environment")(self.h, _p(rgba), w, h, int(bool(use)), int(bool(auto)), _p(col), C.c_float(intensity))

    def set_environment_f32(self, rgb, use=True, auto=False, color=(0, 0, 0), intensity=1.0):
        """float RGB map (h, w, 3): the superset of set_environment behind b200pt_set_environment_f32"""
        rgb = np.ascontiguousarray(rgb, np.float32)
        col = np.asarray(color, np.float32)
        self._f("set_environment_f32")(self.h, _p(rgb), rgb.shape[1], rgb.shape[0], int(bool(use)), int(bool(auto)), _p(col),
                                       C.c_float(intensity))
        return self

    def set_camera(self, origin, llc, horizontal, vertical):
        a = [np.ascontiguousarray(x, np.float32) for x in (origin, llc, horizontal, vertical)]
        self._f("set_camera")(self.h, *[_p(x) for x in a])

    def set_render(self, w, h, spp, depth):
        self.width, self.height, self.spp, self.depth = int(w), int(h), int(spp), int(depth)
        self._f("set_render")(self.h, self.width, self.height, self.spp, self.depth)

    def close(self):
        if getattr(self, "h", None):
            self._f("destroy")(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class Oracle(_SceneChecker):
    prefix = "oracle_"

    def __init__(self, det=False):
        ensure_built()
        self.L = C.CDLL(str(REPO / ("oracle/liboracle_det.so" if det else "oracle/liboracle.so")))
        L = self.L
        P = C.c_void_p
        L.oracle_create.restype = P
        L.oracle_destroy.argtypes = [P]
        L.oracle_add_mesh.argtypes = [P, P, P, P, C.c_int, P, C.c_int, C.c_int, P, C.c_int, C.c_int]
        L.oracle_set_materials.argtypes = [P, P, C.c_int]
        L.oracle_set_environment.argtypes = [P, P, C.c_int, C.c_int, C.c_int, C.c_int, P, C.c_float]
        L.oracle_set_environment_f32.argtypes = [P, P, C.c_int, C.c_int, C.c_int, C.c_int, P, C.c_float]
        L.oracle_set_camera.argtypes = [P, P, P, P, P]
        L.oracle_set_render.argtypes = [P, C.c_int, C.c_int, C.c_int, C.c_int]
        L.oracle_build.argtypes = [P]
        L.oracle_set_nee.argtypes = [P, C.c_int]
        L.oracle_set_output.argtypes = [P, C.c_int]
        L.oracle_shadow_rays.argtypes = [P]
        L.oracle_shadow_rays.restype = C.c_uint64
        L.oracle_camera.argtypes = [P, P, P, C.c_float, C.c_int, C.c_int, P]
        L.oracle_render.argtypes = [P] + [C.c_int] * 9 + [P, P, P]
        L.oracle_render.restype = C.c_uint64
        L.oracle_primary_hits.argtypes = [P, C.c_int, C.c_int, P]
        L.oracle_trace.argtypes = [P, P, P, C.c_int, C.c_float, C.c_float, C.c_int, P]
        L.oracle_sample_disney.argtypes = [P, C.c_int] + [P] * 9
        L.oracle_rng.argtypes = [C.c_uint32, C.c_uint32, C.c_int, P, P]
        L.oracle_math.argtypes = [C.c_int, C.c_int, P, P, P]
        L.oracle_quantise.argtypes = [C.c_float] * 3
        L.oracle_quantise.restype = C.c_uint32
        L.oracle_onb.argtypes = [P, P, C.c_int, P]
        L.oracle_pdf_area_to_solid_angle.argtypes = [C.c_float] * 3
        L.oracle_pdf_area_to_solid_angle.restype = C.c_float
        L.oracle_power_heuristic.argtypes = [C.c_float] * 2
        L.oracle_power_heuristic.restype = C.c_float
        self.det = bool(L.oracle_uses_detmath())
        assert self.det == det
        self.h = L.oracle_create()

    def build(self):
        self.L.oracle_build(self.h)

    def render(self, rank=0, world=1, shard_mode=0, window=None, brute=False, threads=NCPU):
        """Returns dict(sum=(H,W,4) un-normalised, index [y][x] as launched; mean=(H,W,4) flipped;
        rgba=(H,W) flipped; rays)."""
        W, H = self.width, self.height
        x0, y0, x1, y1 = window or (0, 0, W, H)
        s = np.zeros((H, W, 4), np.float32)
        mean = np.zeros((H, W, 4), np.float32)
        rgba = np.zeros((H, W), np.uint32)
        rays = self.L.oracle_render(self.h, rank, world, shard_mode, x0, y0, x1, y1, int(brute), threads, _p(s),
                                    _p(mean), _p(rgba))
        return dict(sum=s, mean=mean, rgba=rgba, rays=int(rays), shadow_rays=int(self.L.oracle_shadow_rays(self.h)))

    def set_nee(self, on=True):
        """Next-event estimation mode (light sampling + MIS); off = the reference's estimator."""
        self.L.oracle_set_nee(self.h, int(on))
        return self

    def set_output(self, transfer):
        """0: linear RGBA8 (reference as shipped); 1: saturate(pow(c, 1/2.2)), device.cu:248."""
        self.L.oracle_set_output(self.h, int(transfer))
        return self

    def primary_hits(self, brute=False, threads=NCPU):
        out = np.empty((self.height, self.width), HIT_DTYPE)
        self.L.oracle_primary_hits(self.h, int(brute), threads, _p(out))
        return out

    def trace(self, o, d, tmin=1e-3, tmax=1e10, brute=False):
        o = np.ascontiguousarray(o, np.float32)
        d = np.ascontiguousarray(d, np.float32)
        out = np.empty(o.shape[0], HIT_DTYPE)
        self.L.oracle_trace(self.h, _p(o), _p(d), o.shape[0], tmin, tmax, int(brute), _p(out))
        return out

    def sample_disney(self, mats, material, wo, lobe_in, rng_in):
        return _sample_disney(self.L.oracle_sample_disney, mats, material, wo, lobe_in, rng_in)

    def rng(self, x, y, n):
        out = np.empty(n, np.float32)
        st = C.c_uint32()
        self.L.oracle_rng(x, y, n, _p(out), C.byref(st))
        return out, st.value

    def math(self, fn, x, y=None):
        return _math(self.L.oracle_math, fn, x, y)

    def onb(self, n, w):
        n = np.ascontiguousarray(n, np.float32)
        w = np.ascontiguousarray(w, np.float32)
        out = np.empty((n.shape[0], 4, 3), np.float32)
        self.L.oracle_onb(_p(n), _p(w), n.shape[0], _p(out))
        return dict(t=out[:, 0], b=out[:, 1], local=out[:, 2], world=out[:, 3])

    def camera(self, look_from, look_at, up, vfov, w, h):
        a = [np.asarray(v, np.float32) for v in (look_from, look_at, up)]
        out = np.empty(12, np.float32)
        self.L.oracle_camera(_p(a[0]), _p(a[1]), _p(a[2]), vfov, w, h, _p(out))
        return out.reshape(4, 3)


FN = dict(sin=0, cos=1, tan=2, atan=3, asin=4, log=5, atan2=6, pow=7)


def _math(f, fn, x, y):
    x = np.ascontiguousarray(x, np.float32)
    y = np.ascontiguousarray(x if y is None else y, np.float32)
    out = np.empty_like(x)
    f(FN[fn], x.size, _p(x), _p(y), _p(out))
    return out


def _sample_disney(f, mats, material, wo, lobe_in, rng_in):
    mats = np.ascontiguousarray(np.ascontiguousarray(mats).view(np.float32).reshape(-1, 17))
    material = np.ascontiguousarray(material, np.int32)
    wo = np.ascontiguousarray(wo, np.float32)
    lobe_in = np.ascontiguousarray(lobe_in, np.int32)
    rng_in = np.ascontiguousarray(rng_in, np.uint32)
    n = material.shape[0]
    fo, wi = np.empty((n, 3), np.float32), np.empty((n, 3), np.float32)
    pdf, lobe, rng = np.empty(n, np.float32), np.empty(n, np.int32), np.empty(n, np.uint32)
    f(_p(mats), n, _p(material), _p(wo), _p(lobe_in), _p(rng_in), _p(fo), _p(wi), _p(pdf), _p(lobe), _p(rng))
    return dict(f=fo, wi=wi, pdf=pdf, lobe=lobe, rng=rng)


def unpack_pairs(blocks):
    """(n_blocks, 20) float32 TriPair blocks (pt_bvh8.h) -> per-triangle arrays of the non-empty halves,
    in block order: ids (uint32), v0, e1, e2 (float32 x 3), block index."""
    b = np.ascontiguousarray(blocks, np.float32).reshape(-1, 20)
    ids = b[:, 18:20].view(np.uint32)                         # (n, 2)
    comp = b[:, :18].reshape(-1, 9, 2)                        # component k of triangle A/B
    tri = np.transpose(comp, (0, 2, 1)).reshape(-1, 9)        # (2n, 9): A0, B0, A1, B1, ...
    idf = ids.reshape(-1)
    keep = idf != 0xFFFFFFFF
    blk = np.repeat(np.arange(b.shape[0]), 2)
    return idf[keep], tri[keep, 0:3], tri[keep, 3:6], tri[keep, 6:9], blk[keep]


class Emu(_SceneChecker):
    prefix = "emu_"

    def __init__(self):
        ensure_built()
        self.L = C.CDLL(str(REPO / "tests/emu/libemu.so"))
        L = self.L
        P = C.c_void_p
        L.emu_create.restype = P
        L.emu_destroy.argtypes = [P]
        L.emu_add_mesh.argtypes = [P, P, P, P, C.c_int, P, C.c_int, C.c_int, P, C.c_int, C.c_int]
        L.emu_set_materials.argtypes = [P, P, C.c_int]
        L.emu_set_environment.argtypes = [P, P, C.c_int, C.c_int, C.c_int, C.c_int, P, C.c_float]
        L.emu_set_environment_f32.argtypes = [P, P, C.c_int, C.c_int, C.c_int, C.c_int, P, C.c_float]
        L.emu_set_camera.argtypes = [P, P, P, P, P]
        L.emu_set_render.argtypes = [P, C.c_int, C.c_int, C.c_int, C.c_int]
        L.emu_build.argtypes = [P]
        L.emu_levels.argtypes = [P]
        L.emu_reinsert_stats.argtypes = [P, C.POINTER(C.c_double), C.POINTER(C.c_double)]
        L.emu_reinsert_stats.restype = C.c_long
        L.emu_update_mesh.argtypes = [P, C.c_int, P, P, C.c_int]
        L.emu_refit.argtypes = [P]
        L.emu_walk_counters.argtypes = [P, P]
        L.emu_force_lbvh.argtypes = [P, C.c_int]
        L.emu_set_nee.argtypes = [P, C.c_int]
        L.emu_shadow_rays.argtypes = [P]
        L.emu_shadow_rays.restype = C.c_uint64
        L.emu_export.argtypes = [P, P, P]
        L.emu_export.restype = C.c_uint32
        L.emu_n_pairs.argtypes = [P]
        L.emu_n_pairs.restype = C.c_uint32
        L.emu_import.argtypes = [P, P, C.c_uint32, P, C.c_uint32]
        L.emu_trace.argtypes = [P, P, P, C.c_int, C.c_float, C.c_float, P, P]
        L.emu_primary_hits.argtypes = [P, C.c_int, P]
        L.emu_render.argtypes = [P, C.c_int, C.c_int, C.c_int, C.c_int, P, P, P]
        L.emu_render.restype = C.c_uint64
        L.emu_sample_disney.argtypes = [P, C.c_int] + [P] * 9
        L.emu_math.argtypes = [C.c_int, C.c_int, P, P, P]
        self.h = L.emu_create()
        self.n_nodes = 0

    def build(self):
        r = self.L.emu_build(self.h)
        assert r > 0, f"emu_build failed: {r}"
        self.n_nodes = r
        return r

    def update_mesh(self, mesh, verts, normals=None):
        verts = np.ascontiguousarray(verts, np.float32)
        normals = None if normals is None else np.ascontiguousarray(normals, np.float32)
        assert self.L.emu_update_mesh(self.h, int(mesh), _p(verts), None if normals is None else _p(normals), verts.shape[0]) == 0

    def refit(self):
        assert self.L.emu_refit(self.h) == 0

    @property
    def levels(self):
        return self.L.emu_levels(self.h)

    def reinsert_stats(self):
        """(moves applied, binary-tree inner-node area after clustering, after the reinsertion passes)"""
        a, b = C.c_double(), C.c_double()
        moves = self.L.emu_reinsert_stats(self.h, C.byref(a), C.byref(b))
        return int(moves), a.value, b.value

    def import_bvh(self, nodes, tris):
        nodes = np.ascontiguousarray(nodes)
        tris = np.ascontiguousarray(tris)
        self.L.emu_import(self.h, _p(nodes), nodes.shape[0], _p(tris), tris.shape[0])

    def render(self, rank=0, world=1, shard_mode=0, threads=NCPU):
        W, H = self.width, self.height
        s = np.zeros((H, W, 4), np.float32)
        mean = np.zeros((H, W, 4), np.float32)
        rgba = np.zeros((H, W), np.uint32)
        rays = self.L.emu_render(self.h, rank, world, shard_mode, threads, _p(s), _p(mean), _p(rgba))
        wc = np.zeros(2, np.uint64)
        self.L.emu_walk_counters(self.h, _p(wc))
        return dict(sum=s, mean=mean, rgba=rgba, rays=int(rays), nodes_visited=int(wc[0]), tris_tested=int(wc[1]),
                    shadow_rays=int(self.L.emu_shadow_rays(self.h)))

    def set_nee(self, on=True):
        self.L.emu_set_nee(self.h, int(on))
        return self

    def export(self):
        """nodes (n, 80) uint8 and triangle-pair blocks (n_blocks, 20) float32 of the emulated build"""
        nodes = np.zeros((self.n_nodes, 80), np.uint8)
        tris = np.zeros((int(self.L.emu_n_pairs(self.h)), 20), np.float32)
        self.L.emu_export(self.h, nodes.ctypes.data, tris.ctypes.data)
        return nodes, tris

    def primary_hits(self, threads=NCPU):
        out = np.empty((self.height, self.width), HIT_DTYPE)
        self.L.emu_primary_hits(self.h, threads, _p(out))
        return out

    def trace(self, o, d, tmin=1e-3, tmax=1e10):
        o = np.ascontiguousarray(o, np.float32)
        d = np.ascontiguousarray(d, np.float32)
        out = np.empty(o.shape[0], HIT_DTYPE)
        cnt = np.zeros(2, np.uint64)
        self.L.emu_trace(self.h, _p(o), _p(d), o.shape[0], tmin, tmax, _p(out), _p(cnt))
        self.last_counters = cnt
        return out

    def sample_disney(self, mats, material, wo, lobe_in, rng_in):
        return _sample_disney(self.L.emu_sample_disney, mats, material, wo, lobe_in, rng_in)

    def math(self, fn, x, y=None):
        return _math(self.L.emu_math, fn, x, y)


def rel_mse(img, ref, eps=1e-2):
    """relMSE = mean_px(|I-R|^2 / (|R|^2 + eps)) on RGB."""
    a = np.asarray(img, np.float64)[..., :3]
    b = np.asarray(ref, np.float64)[..., :3]
    return float(np.mean(np.sum((a - b) ** 2, -1) / (np.sum(b * b, -1) + eps)))


def rel_variance(sum4, n):
    """b200pt_pass_info::rel_variance restated in numpy (binary64): mean over pixels of the variance of the pixel
    mean relative to mean^2 + 1e-2, in Rec.709 luminance, from sum.xyz (sample sum) and sum.w (sum of squared
    sample luminances) after n samples."""
    s = np.asarray(sum4, np.float64).reshape(-1, 4)
    if n < 2:
        return 0.0
    m = (0.2126 * s[:, 0] + 0.7152 * s[:, 1] + 0.0722 * s[:, 2]) / n
    s2 = np.maximum(0.0, s[:, 3] / n - m * m) * (n / (n - 1.0))
    rel = (s2 / n) / (m * m + 1e-2)
    rel = np.where(np.isnan(rel), 0.0, rel)
    return float(rel.sum() / rel.size)


# --------------------------------------------------------------------------------------------------
# The reference itself, compiled for the host (oracle/ref -> oracle/_ref/*.so).  Built here, where
# /root/reference exists; the .so files travel to the GPU box.  When neither exists the tests that
# need them are skipped (the committed golden vectors generated FROM them still run).
REF_DIR = REPO / "oracle" / "_ref"


def ref_available():
    if Path("/root/reference/path_tracer/src/device/device.cu").exists():
        subprocess.run(["make", "-s", "-C", str(REPO / "oracle" / "ref")], check=True)
    return (REF_DIR / "libref_device.so").exists() and (REF_DIR / "libref_host.so").exists()


class Ref(_SceneChecker):
    """/root/reference/path_tracer/src/device/device.cu + disney/*.cuh + math/random/sample_methods/camera, run on the CPU.
    The closest hit (closed-source OptiX in the reference) comes from an Oracle passed to ``render``."""
    prefix = "ref_"

    def __init__(self):
        assert ref_available(), "oracle/_ref is not built"
        self.L = C.CDLL(str(REF_DIR / "libref_device.so"))
        L = self.L
        P = C.c_void_p
        L.ref_create.restype = P
        L.ref_destroy.argtypes = [P]
        L.ref_add_mesh.argtypes = [P, P, P, P, C.c_int, P, C.c_int, C.c_int, P, C.c_int, C.c_int]
        L.ref_set_materials.argtypes = [P, P, C.c_int]
        L.ref_set_environment.argtypes = [P, P, C.c_int, C.c_int, C.c_int, C.c_int, P, C.c_float]
        L.ref_set_camera.argtypes = [P, P, P, P, P]
        L.ref_set_render.argtypes = [P, C.c_int, C.c_int, C.c_int, C.c_int]
        L.ref_set_tracer.argtypes = [P, P, P]
        L.ref_render.argtypes = [P] + [C.c_int] * 5 + [P, P]
        L.ref_render.restype = C.c_uint64
        L.ref_sample_disney.argtypes = [P, C.c_int] + [P] * 9
        L.ref_rng.argtypes = [C.c_uint32, C.c_uint32, C.c_int, P, P]
        L.ref_camera.argtypes = [P, P, P, C.c_float, C.c_int, C.c_int, P]
        L.ref_onb.argtypes = [P, P, C.c_int, P]
        L.ref_uv_on_sphere.argtypes = [P, C.c_int, P]
        L.ref_quantise.argtypes = [C.c_float] * 3
        L.ref_quantise.restype = C.c_uint32
        L.ref_triangle_area.argtypes = [P, P, P]
        L.ref_triangle_area.restype = C.c_float
        L.ref_pdf_area_to_solid_angle.argtypes = [C.c_float] * 3
        L.ref_pdf_area_to_solid_angle.restype = C.c_float
        L.ref_power_heuristic.argtypes = [C.c_float] * 4
        L.ref_power_heuristic.restype = C.c_float
        L.ref_uniform_sample_triangle.argtypes = [C.c_float, C.c_float, P]
        self.h = L.ref_create()

    def build(self):
        pass   # no acceleration structure of its own: see render(tracer=...)

    def render(self, tracer: "Oracle", window=None, threads=NCPU):
        """One launch of the reference's __raygen__ray_gen per pixel; closest hits by ``tracer`` (an Oracle holding the
        same scene).  Returns mean (H, W, 3) float = the argument of make_rgba, rgba (H, W), rays."""
        W, H = self.width, self.height
        x0, y0, x1, y1 = window or (0, 0, W, H)
        self._tracer = tracer   # keep alive
        fn = C.cast(tracer.L.oracle_trace, C.c_void_p)
        self.L.ref_set_tracer(self.h, fn, tracer.h)
        mean = np.zeros((H, W, 3), np.float32)
        rgba = np.zeros((H, W), np.uint32)
        rays = self.L.ref_render(self.h, x0, y0, x1, y1, threads, _p(mean), _p(rgba))
        return dict(mean=mean, rgba=rgba, rays=int(rays))

    def sample_disney(self, mats, material, wo, lobe_in, rng_in):
        return _sample_disney(self.L.ref_sample_disney, mats, material, wo, lobe_in, rng_in)

    def rng(self, x, y, n):
        out = np.empty(n, np.float32)
        st = C.c_uint32()
        self.L.ref_rng(x, y, n, _p(out), C.byref(st))
        return out, st.value

    def camera(self, look_from, look_at, up, vfov, w, h):
        a = [np.asarray(v, np.float32) for v in (look_from, look_at, up)]
        out = np.empty(12, np.float32)
        self.L.ref_camera(_p(a[0]), _p(a[1]), _p(a[2]), vfov, w, h, _p(out))
        return out.reshape(4, 3)

    def onb(self, n, w):
        n = np.ascontiguousarray(n, np.float32)
        w = np.ascontiguousarray(w, np.float32)
        out = np.empty((n.shape[0], 4, 3), np.float32)
        self.L.ref_onb(_p(n), _p(w), n.shape[0], _p(out))
        return dict(t=out[:, 0], b=out[:, 1], local=out[:, 2], world=out[:, 3])

    def uv_on_sphere(self, d):
        d = np.ascontiguousarray(d, np.float32)
        out = np.empty((d.shape[0], 2), np.float32)
        self.L.ref_uv_on_sphere(_p(d), d.shape[0], _p(out))
        return out
