"""create_mesh's vertex de-duplication on the device (b200pt_dedup_mesh, csrc/mesh_dedup.cu) against the host
routine and against the reference's own create_mesh (utils/mesh_loader.cpp:9-83 on tinyobjloader, oracle/_ref) — byte for
byte: local vertex ids in first-occurrence order, first-occurrence normals / texcoords (Q17)."""
import ctypes as C
import subprocess

import numpy as np
import pytest

import owl_path_tracer_b200 as pt
from checkers import REF_DIR, REPO, ref_available
from test_ref_host import assert_same_shapes, our_shapes, synthetic_obj

pytestmark = pytest.mark.gpu
P = C.c_void_p


def gpu_shapes(ctx, path, threads=0):
    H = pt.host_lib()
    h = P()
    if H.b200pt_obj_load_gpu(ctx._h, str(path).encode(), threads, C.byref(h)) != 0:
        raise RuntimeError(H.b200pt_host_last_error().decode())
    out = []
    for i in range(H.b200pt_obj_count(h)):
        v = pt.EntityView()
        assert H.b200pt_obj_shape(h, i, C.byref(v)) == 0
        nn, ntc = v.tex_w, v.tex_h
        out.append((v.name.decode(), pt._view(v.idx, (v.nt, 3), np.int32).copy(), pt._view(v.verts, (v.nv, 3), np.float32).copy(),
                    pt._view(v.normals, (nn, 3), np.float32).copy() if nn else np.zeros((0, 3), np.float32),
                    pt._view(v.uvs, (ntc, 2), np.float32).copy() if ntc else np.zeros((0, 2), np.float32)))
    H.b200pt_obj_free(h)
    return out


def test_device_dedup_equals_host_and_reference(tmp_path):
    ctx = pt.Context()
    files = [REPO / "assets" / "cornell-box.obj.scene", REPO / "assets" / "cube.obj.scene"]
    for name, kw in (("sphere", dict(env="auto")), ("mitsuba", dict(env="auto", detail=2)), ("dragon", dict(detail=60000))):
        sc = pt.Scene(scene=name, width=8, height=8, spp=1, depth=1, **kw)
        f = tmp_path / f"{name}.obj.scene"
        sc.write_obj(f)
        files.append(f)
    for seed in (5, 6):    # groups, relative indices, quads (fan-split), corners without a normal index (irregular: host routine)
        f = tmp_path / f"syn{seed}.obj.scene"
        f.write_text(synthetic_obj(seed, quads=True), newline="")
        files.append(f)
    ref = None
    if ref_available():
        from test_ref_host import ref_shapes
        R = C.CDLL(str(REF_DIR / "libref_host.so"))
        R.ref_host_last_error.restype = C.c_char_p
        R.ref_obj_load.restype = P; R.ref_obj_load.argtypes = [C.c_char_p]
        R.ref_obj_free.argtypes = [P]; R.ref_obj_count.argtypes = [P]
        R.ref_obj_name.argtypes = [P, C.c_int]; R.ref_obj_name.restype = C.c_char_p
        R.ref_obj_sizes.argtypes = [P, C.c_int, P]; R.ref_obj_copy.argtypes = [P, C.c_int, P, P, P, P]
        ref = lambda p: ref_shapes(R, p)
    total = 0
    for f in files:
        host = our_shapes(f)
        dev = gpu_shapes(ctx, f)
        assert_same_shapes(dev, host)
        total += sum(s[1].shape[0] for s in dev)
        if ref is not None and "syn" not in f.name:      # quads: the reference mis-indexes them (test_ref_host), triangles only here
            assert_same_shapes(dev, ref(f))
    assert total > 100000
    ctx.close()


def test_dedup_entry_point_directly_and_its_errors():
    ctx = pt.Context()
    L = pt.lib()
    av = np.array([[0, 0, 0], [1, 0, 0], [0, 1, 0], [1, 1, 0], [9, 9, 9]], np.float32)
    an = np.array([[0, 0, 1], [0, 1, 0]], np.float32)
    at = np.array([[0, 0], [1, 1]], np.float32)
    cv = np.array([3, 1, 2, 1, 0, 2, 3, 2, 1], np.int32)          # first occurrences: 3, 1, 2, 0 -> local ids 0, 1, 2, 3
    cn = np.array([1, 0, 0, 1, 1, 1, 0, 0, 0], np.int32)
    ct = np.array([0, 1, 0, 1, 1, 1, 1, 1, 1], np.int32)
    ov, on, ot = np.zeros((9, 3), np.float32), np.zeros((9, 3), np.float32), np.zeros((9, 2), np.float32)
    oi, nv = np.zeros(9, np.int32), C.c_int32()
    def call(cn_=cn, ct_=ct, cv_=cv):
        return L.b200pt_dedup_mesh(ctx._h, av.ctypes.data, 5, an.ctypes.data, 2, at.ctypes.data, 2, cv_.ctypes.data,
                                   None if cn_ is None else cn_.ctypes.data, None if ct_ is None else ct_.ctypes.data, 9,
                                   ov.ctypes.data, on.ctypes.data, ot.ctypes.data, oi.ctypes.data, C.byref(nv))
    assert call() == 0 and nv.value == 4
    assert oi.tolist() == [0, 1, 2, 1, 3, 2, 0, 2, 1]
    assert ov[:4].tolist() == av[[3, 1, 2, 0]].tolist()
    assert on[:4].tolist() == an[[1, 0, 0, 1]].tolist()           # the normal of each vertex's FIRST corner
    assert ot[:4].tolist() == at[[0, 1, 0, 1]].tolist()
    assert call(cn_=None, ct_=None) == 0 and nv.value == 4
    bad = cn.copy(); bad[4] = -1
    assert call(cn_=bad) != 0 and b"irregular" in L.b200pt_last_error()
    bad = cv.copy(); bad[2] = 5
    assert call(cv_=bad) != 0 and b"out of range" in L.b200pt_last_error()
    ctx.close()


def test_path_tracer_gpu_dedup_writes_the_same_frame(tmp_path):
    exe = REPO / "owl_path_tracer_b200" / "bin" / "path_tracer"
    args = ["--assets", str(REPO / "assets"), "--scene", "cornell-box", "--size", "96", "64", "--spp", "4", "--depth", "5",
            "--env", "color", "--env-intensity", "0"]
    outs = []
    for extra in ([], ["--gpu-dedup"]):
        d = tmp_path / ("g" if extra else "h")
        d.mkdir()
        r = subprocess.run([str(exe), *args, *extra, "--out", str(d)], capture_output=True, text=True, timeout=300)
        assert r.returncode == 0, r.stderr[-500:]
        outs.append({f.name: f.read_bytes() for f in sorted(d.glob("*.png"))})
    assert len(outs[0]) == 5 and outs[0] == outs[1]
