"""The product's host loaders against the REFERENCE'S OWN loaders (oracle/_ref/libref_host.so = utils/mesh_loader.cpp on the
vendored tinyobjloader, utils/image_buffer.cpp on stb_image, utils/parser.cpp on nlohmann json, camera.cpp; see oracle/ref/).
Everything is compared byte for byte.  Skipped where oracle/_ref is not built."""
import ctypes as C
import json
import struct
import zlib

import numpy as np
import pytest

import owl_path_tracer_b200 as pt
from checkers import REF_DIR, REPO, ref_available

pytestmark = pytest.mark.skipif(not ref_available(), reason="oracle/_ref not built (no /root/reference here)")
P = C.c_void_p


class Settings(C.Structure):    # b200pt_settings_view == ref_settings
    _fields_ = [("scene", C.c_char * 256), ("test_name", C.c_char * 256), ("test_material", C.c_char * 256),
                ("test_attribute", C.c_char * 256), ("buffer_size", C.c_int32 * 2), ("max_samples", C.c_int32),
                ("max_path_depth", C.c_int32), ("environment_use", C.c_int32), ("environment_auto", C.c_int32),
                ("test_material_type", C.c_int32), ("environment_color", C.c_float * 3), ("environment_intensity", C.c_float),
                ("test_step_size", C.c_float), ("n_vec_values", C.c_int32), ("n_flt_values", C.c_int32),
                ("vec_values", C.c_float * 3 * 64), ("flt_values", C.c_float * 64)]


@pytest.fixture(scope="module")
def R():
    L = C.CDLL(str(REF_DIR / "libref_host.so"))
    L.ref_host_last_error.restype = C.c_char_p
    L.ref_obj_load.restype = P
    L.ref_obj_load.argtypes = [C.c_char_p]
    L.ref_obj_free.argtypes = [P]
    L.ref_obj_count.argtypes = [P]
    L.ref_obj_name.argtypes = [P, C.c_int]
    L.ref_obj_name.restype = C.c_char_p
    L.ref_obj_sizes.argtypes = [P, C.c_int, P]
    L.ref_obj_copy.argtypes = [P, C.c_int, P, P, P, P]
    L.ref_image_load.restype = P
    L.ref_image_load.argtypes = [C.c_char_p, C.c_char_p, P, P]
    L.ref_image_pixels.restype = P
    L.ref_image_pixels.argtypes = [P]
    L.ref_image_free.argtypes = [P]
    L.ref_image_write.argtypes = [C.c_int32, C.c_int32, P, C.c_char_p, C.c_char_p]
    L.ref_parse_materials.restype = P
    L.ref_parse_materials.argtypes = [C.c_char_p]
    L.ref_materials_free.argtypes = [P]
    L.ref_materials_count.argtypes = [P]
    L.ref_material_name.argtypes = [P, C.c_int]
    L.ref_material_name.restype = C.c_char_p
    L.ref_material_file.argtypes = [P, C.c_int]
    L.ref_material_file.restype = C.c_char_p
    L.ref_material_data.argtypes = [P, C.c_int, P]
    L.ref_parse_camera.argtypes = [C.c_char_p, C.c_int, C.c_int, P]
    L.ref_parse_settings.argtypes = [C.c_char_p, P]
    return L


def ref_shapes(R, path):
    h = R.ref_obj_load(str(path).encode())
    if not h:
        raise RuntimeError(R.ref_host_last_error().decode())
    out = []
    for i in range(R.ref_obj_count(h)):
        sz = np.zeros(4, np.int64)
        R.ref_obj_sizes(h, i, sz.ctypes.data)
        idx, v, n, t = (np.zeros((sz[0], 3), np.int32), np.zeros((sz[1], 3), np.float32), np.zeros((sz[2], 3), np.float32),
                        np.zeros((sz[3], 2), np.float32))
        R.ref_obj_copy(h, i, idx.ctypes.data, v.ctypes.data, n.ctypes.data, t.ctypes.data)
        out.append((R.ref_obj_name(h, i).decode(), idx, v, n, t))
    R.ref_obj_free(h)
    return out


def our_shapes(path, threads=0):
    H = pt.host_lib()
    h = P()
    if H.b200pt_obj_load(str(path).encode(), threads, C.byref(h)) != 0:
        raise RuntimeError(H.b200pt_host_last_error().decode())
    out = []
    for i in range(H.b200pt_obj_count(h)):
        v = pt.EntityView()
        assert H.b200pt_obj_shape(h, i, C.byref(v)) == 0
        nn, ntc = v.tex_w, v.tex_h      # counts of normals / texcoords (see b200pt_obj_shape)
        out.append((v.name.decode(), pt._view(v.idx, (v.nt, 3), np.int32).copy(), pt._view(v.verts, (v.nv, 3), np.float32).copy(),
                    pt._view(v.normals, (nn, 3), np.float32).copy() if nn else np.zeros((0, 3), np.float32),
                    pt._view(v.uvs, (ntc, 2), np.float32).copy() if ntc else np.zeros((0, 2), np.float32)))
    H.b200pt_obj_free(h)
    return out


def assert_same_shapes(a, b):
    assert [x[0] for x in a] == [x[0] for x in b]
    for x, y in zip(a, b):
        for u, v in zip(x[1:], y[1:]):
            assert u.shape == v.shape, (x[0], u.shape, v.shape)
            assert (u.view(np.uint32) == v.view(np.uint32)).all(), x[0]


# ------------------------------------------------------------------ OBJ (tinyobjloader + create_mesh)

@pytest.mark.parametrize("name", ["cornell-box", "cube"])
def test_bundled_obj_files(R, name):
    f = REPO / "assets" / f"{name}.obj.scene"
    ref = ref_shapes(R, f)
    assert len(ref) > 0
    for threads in (0, 1, 3, 8):
        assert_same_shapes(our_shapes(f, threads), ref)


def test_generated_stand_ins_round_trip_through_both_loaders(R, tmp_path):
    """every procedural stand-in, written as .obj.scene (b200pt_write_obj) and read back by both loaders"""
    for name, kw in (("sphere", dict(env="auto")), ("mitsuba", dict(env="auto", detail=1)), ("car", dict(detail=1)),
                     ("dragon", dict(detail=1))):
        sc = pt.Scene(scene=name, width=8, height=8, spp=1, depth=1, **kw)
        f = tmp_path / f"{name}.obj.scene"
        sc.write_obj(f)
        ref = ref_shapes(R, f)
        assert sum(s[1].shape[0] for s in ref) == sc.n_triangles
        assert_same_shapes(our_shapes(f), ref)
        assert_same_shapes(our_shapes(f, 5), ref)


def synthetic_obj(seed, quads=False):
    rng = np.random.default_rng(seed)
    lines, nv = ["# synthetic", "mtllib nothing.mtl"], 0
    for sidx in range(12):
        if sidx % 3 == 0:
            lines.append(f"o m{sidx}" + ("  " if sidx % 2 else ""))        # trailing blanks are part of an `o` name
        elif sidx % 3 == 1:
            lines.append(f"g  m{sidx}\t extra words ")                     # `g`: words joined by one blank
        else:
            lines.append(f"   g m{sidx}")
        for _ in range(int(rng.integers(40, 300))):
            k = int(rng.integers(3, 6))
            for _ in range(k):
                x, y, z = rng.normal(size=3)
                fmt = ["{:.6f}", "{:.9g}", "{:e}", "{:.3f}", "{:+.4f}"][int(rng.integers(0, 5))]
                lines.append("v " + " ".join(fmt.format(c) for c in (x * 10 ** int(rng.integers(-3, 4)), y, z)) +
                             (" 0.5 0.25 1" if rng.random() < 0.1 else ""))        # vertex colours are ignored
                lines.append(f"vn {x:.3f} {y:.3f} {z:.3f}")
                lines.append(f"vt {abs(x) % 1:.4f} {abs(y) % 1:.4f}" + (" 0" if rng.random() < 0.2 else ""))
            nv += k
            if rng.random() < 0.2:
                lines.append("# a comment between attributes and faces")
            if rng.random() < 0.1:
                lines.append("usemtl whatever")
            if rng.random() < 0.1:
                lines.append("s 1")
            for _ in range(int(rng.integers(1, 4))):
                nc = 3 if not quads else int(rng.integers(3, min(k, 4) + 1))
                corners = []
                for c in rng.permutation(k)[:nc]:
                    if rng.random() < 0.5:   # relative: -1 is the attribute defined last
                        a = -(k - int(c))
                        corners.append(f"{a}/{a}/{a}" if rng.random() < 0.7 else f"{a}//{a}")
                    else:
                        a = nv - k + int(c) + 1
                        corners.append([f"{a}/{a}/{a}", f"{a}//{a}", f"{a}/{a}"][int(rng.integers(0, 3))])
                lines.append("f " + ("  " if rng.random() < 0.3 else " ").join(corners) + (" " if rng.random() < 0.2 else ""))
            if rng.random() < 0.05:
                lines.append(f"f {nv} {nv - 1}")                                    # degenerate face: dropped
    ends = ["\n", "\r\n"]
    return "".join(l + ends[i % 7 == 0] for i, l in enumerate(lines))


@pytest.mark.parametrize("seed", [5, 6, 7])
def test_synthetic_obj_groups_relative_indices_crlf_number_formats(R, tmp_path, seed):
    f = tmp_path / "t.obj.scene"
    f.write_text(synthetic_obj(seed), newline="")
    ref = ref_shapes(R, f)
    assert len(ref) == 12 and ref[1][0] == "m1 extra words" and ref[6][0] == "m6" and ref[3][0] == "m3  "
    for threads in (1, 2, 7, 16):
        assert_same_shapes(our_shapes(f, threads), ref)


def test_number_parsing_follows_tinyobj_not_strtod(R, tmp_path):
    """tinyobj's decimal routine is not correctly rounded; the same digits must give the same float in both loaders."""
    rs = np.random.RandomState(3)
    vals = []
    for _ in range(30000):
        nd = rs.randint(1, 18)
        digits = "".join(str(d) for d in rs.randint(0, 10, nd))
        s = ("-" if rs.rand() < 0.5 else "") + str(rs.randint(0, 1000)) + "." + digits
        if rs.rand() < 0.3:
            s += "e" + str(rs.randint(-30, 30))
        vals.append(s)
    vals += ["1e5", ".5", "-.25", "+3.", "7", "1e", "abc", "1.5e+2", "0.1e-3", "1e400", "-0", "123456789012345678901234567890"]
    lines = ["o s"]
    while len(vals) % 3:
        vals.append("0")
    for i in range(0, len(vals), 3):
        lines.append("v " + " ".join(vals[i:i + 3]))
        lines.append("vn 0 0 1")
    n = len(vals) // 3
    for i in range(1, n - 1, 3):
        lines.append(f"f {i}//{i} {i + 1}//{i + 1} {i + 2}//{i + 2}")
    f = tmp_path / "n.obj.scene"
    f.write_text("\n".join(lines) + "\n")
    ref = ref_shapes(R, f)
    ours = our_shapes(f, 4)
    assert_same_shapes(ours, ref)
    strtod = np.array([float(v) if v not in ("1e", "abc") else 0.0 for v in vals[:3000]], np.float64).astype(np.float32)
    got = ref[0][2].reshape(-1)
    assert got.size > 3000


def test_malformed_face_lines_fail_like_tinyobj(R, tmp_path):
    """`f 1//1 2//1 3//1 # tri` used to loop forever.  tinyobj rejects the file ("Failed parse `f' line") and
    mesh_loader.cpp:99-101 throws; index 0 likewise.  Attribute indices past the end: tinyobj only warns and the reference
    then reads out of bounds — here they are refused."""
    head = "o a\nv 0 0 0\nv 1 0 0\nv 0 1 0\nvn 0 0 1\nvt 0 0\n"
    for bad in ("f 1//1 2//1 3//1 # tri\n", "f 1//1 2//1 x\n", "f 0//1 1//1 2//1\n", "f 1/0/1 2/1/1 3/1/1\n", "f 1//1 2//1 3//\n"):
        f = tmp_path / "bad.obj.scene"
        f.write_text(head + "f 1//1 2//1 3//1\n" + bad)
        with pytest.raises(RuntimeError, match="Failed parse `f' line.*line 8"):
            ref_shapes(R, f)
        for threads in (1, 3):
            with pytest.raises(RuntimeError, match="Failed parse `f' line.*line 8"):
                our_shapes(f, threads)
    for oob, what in (("f 1//1 2//1 3//9\n", "normal"), ("f 1/7/1 2/1/1 3/1/1\n", "texcoord"), ("f 1//1 2//1 9//1\n", "vertex"),
                      ("f 1//1 2//1 -7//1\n", "vertex")):
        f = tmp_path / "oob.obj.scene"
        f.write_text(head + oob)
        with pytest.raises(RuntimeError, match=f"{what} index out of range"):
            our_shapes(f)


def test_faces_with_more_than_three_corners_are_a_documented_difference(R, tmp_path):
    """triangulate=false + `index_offset += 3` (mesh_loader.cpp:27,78): the reference reads a quad as one triangle and then
    starts the next face in the middle of it.  The product fan-splits instead; triangle-only shapes are identical."""
    f = tmp_path / "q.obj.scene"
    f.write_text("o q\nv 0 0 0\nv 1 0 0\nv 1 1 0\nv 0 1 0\nvn 0 0 1\nf 1//1 2//1 3//1 4//1\nf 1//1 3//1 4//1\n")
    ref, ours = ref_shapes(R, f), our_shapes(f)
    assert ref[0][1].tolist() == [[0, 1, 2], [3, 0, 2]]            # second "triangle" = corner 4 of the quad + 2 of the next face
    assert ours[0][1].tolist() == [[0, 1, 2], [0, 2, 3], [0, 2, 3]]


# ------------------------------------------------------------------ images (stb_image)

def write_png(path, img, ctype, depth=8, palette=None, trns=None, interlace=0, filters=None):
    h, w = img.shape[:2]
    ch = {0: 1, 2: 3, 3: 1, 4: 2, 6: 4}[ctype]
    a = img.reshape(h, w, ch)
    def pack_rows(rows):
        out = b""
        for y, row in enumerate(rows):
            if depth == 16:
                raw = row.astype(">u2").tobytes()
            elif depth == 8:
                raw = row.astype(np.uint8).tobytes()
            else:
                bits = np.unpackbits(row.astype(np.uint8).reshape(-1, 1), axis=1)[:, 8 - depth:].reshape(-1)
                raw = np.packbits(bits).tobytes()
            out += b"\x00" + raw
        return out
    if interlace:
        data = b""
        for x0, y0, dx, dy in ((0, 0, 8, 8), (4, 0, 8, 8), (0, 4, 4, 8), (2, 0, 4, 4), (0, 2, 2, 4), (1, 0, 2, 2), (0, 1, 1, 2)):
            sub = a[y0::dy, x0::dx]
            if sub.size:
                data += pack_rows(list(sub))
    else:
        data = pack_rows(list(a))
    def chunk(t, b):
        return struct.pack(">I", len(b)) + t + b + struct.pack(">I", zlib.crc32(t + b))
    out = b"\x89PNG\r\n\x1a\n" + chunk(b"IHDR", struct.pack(">IIBBBBB", w, h, depth, ctype, 0, 0, interlace))
    if palette is not None:
        out += chunk(b"PLTE", palette.astype(np.uint8).tobytes())
    if trns is not None:
        out += chunk(b"tRNS", trns)
    half = len(data) // 2
    comp = zlib.compress(data, 6)
    out += chunk(b"IDAT", comp[:len(comp) // 2]) + chunk(b"IDAT", comp[len(comp) // 2:]) + chunk(b"IEND", b"")
    path.write_bytes(out)


def ref_image(R, name, path):
    w, h = C.c_int32(), C.c_int32()
    hnd = R.ref_image_load(name.encode(), str(path).encode(), C.byref(w), C.byref(h))
    p = R.ref_image_pixels(hnd)
    img = None if not p else np.ctypeslib.as_array(C.cast(p, C.POINTER(C.c_uint32)), (h.value, w.value)).copy()
    R.ref_image_free(hnd)
    return img


def our_image(name, path):
    H = pt.host_lib()
    w, h, p = C.c_int32(), C.c_int32(), P()
    if H.b200pt_image_load(name.encode(), str(path).encode(), C.byref(w), C.byref(h), C.byref(p)) != 0:
        raise RuntimeError(H.b200pt_host_last_error().decode())
    if not p:
        return None
    img = np.ctypeslib.as_array(C.cast(p, C.POINTER(C.c_uint32)), (h.value, w.value)).copy()
    H.b200pt_image_free(p)
    return img


def test_png_decoding_equals_stb(R, tmp_path):
    rs = np.random.RandomState(1)
    cases = []
    for ctype, ch in ((0, 1), (2, 3), (4, 2), (6, 4)):
        for depth in (8, 16):
            for interlace in (0, 1):
                hi = 256 if depth == 8 else 65536
                cases.append((f"c{ctype}d{depth}i{interlace}", rs.randint(0, hi, (13, 17, ch)), ctype, depth, None, None, interlace))
    for depth in (1, 2, 4):
        cases.append((f"g{depth}", rs.randint(0, 2 ** depth, (9, 21, 1)), 0, depth, None, None, 0))
        cases.append((f"g{depth}i", rs.randint(0, 2 ** depth, (9, 21, 1)), 0, depth, None, None, 1))
        cases.append((f"p{depth}", rs.randint(0, 2 ** depth, (9, 21, 1)), 3, depth, rs.randint(0, 256, (2 ** depth, 3)), bytes([0, 128]), 0))
    cases.append(("p8", rs.randint(0, 256, (11, 19, 1)), 3, 8, rs.randint(0, 256, (256, 3)), bytes(range(0, 200, 3)), 0))
    cases.append(("p8i", rs.randint(0, 256, (11, 19, 1)), 3, 8, rs.randint(0, 256, (256, 3)), None, 1))
    cases.append(("key8", rs.randint(0, 4, (8, 8, 3)), 2, 8, None, struct.pack(">HHH", 1, 2, 3), 0))
    cases.append(("keyg", rs.randint(0, 4, (8, 8, 1)), 0, 8, None, struct.pack(">H", 2), 0))
    for name, img, ctype, depth, pal, trns, il in cases:
        write_png(tmp_path / f"{name}.png", img, ctype, depth, pal, trns, il)
        a, b = our_image(f"{name}.png", tmp_path), ref_image(R, f"{name}.png", tmp_path)
        assert b is not None and a is not None, name
        assert a.shape == b.shape and (a == b).all(), name
    # real-world filters: let zlib-independent filtering be exercised through our own writer (filter 0) and stb's (adaptive)
    img = (rs.randint(0, 2 ** 32, (37, 53), dtype=np.uint64).astype(np.uint32)) | np.uint32(0xFF000000)
    smooth = (np.add.outer(np.arange(37), np.arange(53)) % 256).astype(np.uint32) * np.uint32(0x010203) | np.uint32(0xFF000000)
    for k, im in enumerate((img, smooth)):
        assert R.ref_image_write(53, 37, im.ctypes.data, f"stb{k}.png".encode(), str(tmp_path).encode()) == 0
        a, b = our_image(f"stb{k}.png", tmp_path), ref_image(R, f"stb{k}.png", tmp_path)
        assert (a == b).all() and (a == im[::-1]).all()
        assert pt.host_lib().b200pt_write_png(str(tmp_path / f"own{k}.png").encode(), 53, 37, im.ctypes.data) == 0
        assert (ref_image(R, f"own{k}.png", tmp_path) == im[::-1]).all()      # stb reads what the product writes
    assert our_image("nope.png", tmp_path) is None and ref_image(R, "nope.png", tmp_path) is None


def test_hdr_to_ldr_equals_stb(R, tmp_path):
    """environment.hdr goes through stbi_load(..., STBI_rgb_alpha): RGBE -> float -> pow(x, 1/2.2) * 255 + 0.5 (R7 / Q5)."""
    # (a) the generated stand-in environment (new-style RLE)
    sc = pt.Scene(scene="sphere", width=8, height=8, spp=1, depth=1, env="map", env_float=True)
    f32 = np.ascontiguousarray(sc.environment["rgb_f32"][::-1])            # file order: top row first
    H = pt.host_lib()
    rs = np.random.RandomState(2)
    def rgbe_file(path, rgbe, rle):
        h, w = rgbe.shape[:2]
        with open(path, "wb") as f:
            f.write(b"#?RADIANCE\n# comment\nFORMAT=32-bit_rle_rgbe\nEXPOSURE=1.0\n\n-Y %d +X %d\n" % (h, w))
            for y in range(h):
                if rle:
                    f.write(bytes([2, 2, w >> 8, w & 255]))
                    for k in range(4):
                        x = 0
                        while x < w:
                            if rs.rand() < 0.5:      # a run
                                n = min(w - x, int(rs.randint(1, 128)))
                                val = rgbe[y, x, k]
                                f.write(bytes([128 + n, int(val)]))
                                rgbe[y, x:x + n, k] = val
                            else:
                                n = min(w - x, int(rs.randint(1, 129)))
                                f.write(bytes([n]) + rgbe[y, x:x + n, k].tobytes())
                            x += n
                else:
                    f.write(rgbe[y].tobytes())
    for name, w, h, rle in (("rle.hdr", 64, 16, True), ("flat.hdr", 64, 16, False), ("narrow.hdr", 5, 7, False), ("wide.hdr", 300, 3, True)):
        rgbe = rs.randint(0, 256, (h, w, 4)).astype(np.uint8)
        rgbe[..., 3] = rs.randint(100, 150, (h, w))          # exponents around 1.0: both ends of the 8-bit clamp
        rgbe[0, :4, 3] = 0                                    # zero exponent = black
        if not rle:
            rgbe[..., 0] |= 128                               # a flat file whose first pixel is not (2, 2, <128)
        rgbe_file(tmp_path / name, rgbe, rle)
        a, b = our_image(name, tmp_path), ref_image(R, name, tmp_path)
        assert b is not None and a is not None and a.shape == (h, w), name
        assert (a == b).all(), (name, int((a != b).sum()))


def test_existing_but_undecodable_image_is_an_error(tmp_path):
    """stb also reads JPEG / BMP / TGA / GIF / PSD / PIC / PNM; the product reads PNG and Radiance only and says so instead
    of continuing with a stand-in texture."""
    (tmp_path / "x.jpg").write_bytes(b"\xff\xd8\xff\xe0" + b"\0" * 64)
    with pytest.raises(RuntimeError, match="unsupported image format|could not be decoded"):
        our_image("x.jpg", tmp_path)


# ------------------------------------------------------------------ JSON (nlohmann) + camera

SCENES = ["car", "cornell-box", "cube", "dragon", "mitsuba", "sphere"]


@pytest.mark.parametrize("name", SCENES)
def test_materials_and_camera_of_every_bundled_scene(R, name):
    H = pt.host_lib()
    f = str(REPO / "assets" / f"{name}.json").encode()
    a, b = np.zeros(12, np.float32), np.zeros(12, np.float32)
    for w, h in ((1920, 1080), (1080, 1440), (512, 512), (3, 1000)):
        assert H.b200pt_parse_camera(f, w, h, a.ctypes.data) == 0 and R.ref_parse_camera(f, w, h, b.ctypes.data) == 0
        assert (a.view(np.uint32) == b.view(np.uint32)).all()
    rm = R.ref_parse_materials(f)
    om = P()
    ours_ok = H.b200pt_parse_materials(f, C.byref(om)) == 0
    if not rm:      # assets/sphere.json has no "use_texture": nlohmann throws (parser.cpp:33); the product reads it as false
        assert name == "sphere" and ours_ok
        return
    assert ours_ok and H.b200pt_materials_count(om) == R.ref_materials_count(rm) > 0
    for i in range(R.ref_materials_count(rm)):
        assert H.b200pt_materials_name(om, i) == R.ref_material_name(rm, i)
        assert H.b200pt_materials_file(om, i) == R.ref_material_file(rm, i)
        d = np.zeros(17, np.float32)
        R.ref_material_data(rm, i, d.ctypes.data)
        mine = np.ctypeslib.as_array(C.cast(H.b200pt_materials_data(om, i), C.POINTER(C.c_float)), (17,))
        assert (mine.view(np.uint32) == d.view(np.uint32)).all()
    R.ref_materials_free(rm)
    H.b200pt_materials_free(om)


def test_settings_json(R, tmp_path):
    H = pt.host_lib()
    base = json.loads((REPO / "assets/settings.json").read_text())
    variants = [base]
    v = json.loads(json.dumps(base))
    v["test"]["values"] = [[0.1, 0.2, 0.3], [1, 0, 0.5e-1]]
    v["test"]["step_size"] = 1e-1
    v["environment_color"] = [0.25, 1, 2.5E0]
    v["buffer_size"] = [1920, 1080]
    v["environment_intensity"] = 0.30000001192092896
    variants.append(v)
    for k, cfg in enumerate(variants):
        f = tmp_path / f"s{k}.json"
        f.write_text(json.dumps(cfg, indent=k))
        a, b = Settings(), Settings()
        assert H.b200pt_parse_settings(str(f).encode(), C.byref(a)) == 0
        assert R.ref_parse_settings(str(f).encode(), C.byref(b)) == 0
        assert bytes(a) == bytes(b)
        assert a.scene == cfg["scene"].encode() and a.max_samples == cfg["max_samples"]


def test_json_number_grammar_equals_nlohmann(R, tmp_path):
    """floats of every JSON spelling go through parse_materials: same binary32 as nlohmann's get<float>()"""
    H = pt.host_lib()
    rs = np.random.RandomState(4)
    proto = json.loads((REPO / "assets/cube.json").read_text())["materials"][0]
    proto["use_texture"] = False
    keys = [k for k, v in proto.items() if isinstance(v, (int, float)) and not isinstance(v, bool)]
    mats = []
    for i in range(200):
        m = dict(proto, name=f"m{i}")
        for k in keys:
            r = rs.rand()
            m[k] = "@" + (f"{rs.rand() * 10 ** rs.randint(-8, 8):.{rs.randint(1, 17)}g}" if r < 0.6 else
                          f"{rs.randint(-5, 5)}" if r < 0.8 else f"{rs.rand():.12e}".replace("e", "E"))
        m["base_color"] = ["@" + f"{rs.rand():.9f}" for _ in range(3)]
        mats.append(m)
    text = json.dumps({"camera": json.loads((REPO / "assets/cube.json").read_text())["camera"], "materials": mats})
    import re
    text = re.sub(r'"@([^"]+)"', r"\1", text)       # raw number tokens, spelled exactly as generated
    f = tmp_path / "m.json"
    f.write_text(text)
    rm, om = R.ref_parse_materials(str(f).encode()), P()
    assert rm and H.b200pt_parse_materials(str(f).encode(), C.byref(om)) == 0
    for i in range(200):
        d = np.zeros(17, np.float32)
        R.ref_material_data(rm, i, d.ctypes.data)
        mine = np.ctypeslib.as_array(C.cast(H.b200pt_materials_data(om, i), C.POINTER(C.c_float)), (17,))
        assert (mine.view(np.uint32) == d.view(np.uint32)).all(), i
