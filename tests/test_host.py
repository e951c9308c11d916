"""CPU tests of the host surface (loaders, application stages) and of the drop-in boundary:
the C-ABI libraries load, export every symbol the headers declare, and FAIL LOUDLY without a GPU
(no CPU fallback)."""
import ctypes as C
import re
import struct
import zlib

import numpy as np
import pytest

import owl_path_tracer_b200 as pt
from checkers import REPO


def declared_symbols(header):
    text = (REPO / "include" / header).read_text()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(b200pt_[a-z0-9_]+)\s*\(", text)))


def test_c_abi_exports_every_declared_symbol():
    L, H = pt.lib(), pt.host_lib()
    syms = declared_symbols("b200pt.h")
    assert len(syms) >= 20
    for s in syms:
        assert hasattr(L, s), s
    hs = declared_symbols("b200pt_host.h")
    assert len(hs) >= 14
    for s in hs:
        assert hasattr(H, s), s
    assert L.b200pt_version() == 100


def test_struct_layouts_match_header():
    assert pt.MATERIAL_DTYPE.itemsize == 68 and pt.HIT_DTYPE.itemsize == 20
    assert C.sizeof(pt.Config) == 24          # b200pt_config: six int32


def test_no_cpu_fallback():
    """Without a CUDA device the product refuses to run (on the GPU box this test is a no-op)."""
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    with pytest.raises(pt.B200ptError, match="no CUDA device|CUDA"):
        pt.Context()


def test_settings_and_scene_files_of_the_reference():
    sc = pt.Scene(allow_procedural=True, width=16, height=12, spp=1, depth=1)   # settings.json says "car"
    assert sc.material_names()[:3] == ["BodyGlossBlackMat", "BodyMat", "BodyMat_BK"]
    assert len(sc.material_names()) == 12
    sc2 = pt.Scene(scene="cornell-box")
    assert (sc2.width, sc2.height, sc2.spp, sc2.depth) == (1080, 1440, 12288, 16)   # assets/settings.json:3-5
    ents = sc2.entities()
    assert [e["name"] for e in ents] == ["box", "wall_left", "wall_right", "wall_tbb", "sphere", "light"]
    assert sum(e["idx"].shape[0] for e in ents) == 17974
    names = sc2.material_names()
    assert [names[e["material"]] for e in ents] == [e["name"] for e in ents]
    m = sc2.materials()
    assert m[names.index("light")]["emission"] == 15.0 and m[names.index("box")]["metallic"] == 1.0
    with pytest.raises(pt.B200ptError):
        pt.Scene(scene="dragon", allow_procedural=False)     # missing .obj.scene -> error like the reference
    with pytest.raises(pt.B200ptError):
        pt.Scene(scene="no-such-scene")


def test_obj_loader_quirks(tmp_path):
    """mesh_loader.cpp:9-83: vertices shared by position index; first normal/uv wins; shapes whose
    name has no material are dropped (application.cpp:166-179)."""
    (tmp_path / "settings.json").write_text((REPO / "assets/settings.json").read_text().replace('"car"', '"t"'))
    mat = (REPO / "assets/cube.json").read_text().replace('"cube"', '"a"').replace('"use_texture": true', '"use_texture": false')
    (tmp_path / "t.json").write_text(mat)
    (tmp_path / "t.obj.scene").write_text("""
o a
v 0 0 0
v 1 0 0
v 0 1 0
v 1 1 0
vn 0 0 1
vn 0 1 0
vt 0 0
vt 1 1
f 1/1/1 2/1/1 3/1/1
f 2/2/2 4/2/2 3/2/2
f -4//1 -3//1 -1//1
o unbound
v 5 5 5
v 6 5 5
v 5 6 5
vn 1 0 0
f 5//3 6//3 7//3
""")
    sc = pt.Scene(assets=tmp_path, width=8, height=8, spp=1, depth=1)
    ents = sc.entities()
    assert len(ents) == 1 and ents[0]["name"] == "a"
    e = ents[0]
    assert e["verts"].shape == (4, 3) and e["idx"].tolist() == [[0, 1, 2], [1, 3, 2], [0, 1, 3]]
    # vertices 0..2 got normal 1 (first face), vertex 3 got normal 2 when it first appeared
    assert e["normals"].tolist() == [[0, 0, 1]] * 3 + [[0, 1, 0]]
    assert e["uvs"].tolist() == [[0, 0]] * 3 + [[1, 1]]


def test_parallel_obj_parse_equals_the_serial_parse(tmp_path, monkeypatch):
    """the OBJ file is parsed in chunks by concurrent threads and the shapes' vertex maps are built
    concurrently; one chunk / one thread is the serial parser.  Same entities, bit for bit, whatever the thread count —
    on the reference's own cornell-box file and on a synthetic file whose negative (relative) indices, groups, quads,
    CRLF line ends and comments straddle the chunk cuts."""
    import json
    def entities(assets, threads):
        monkeypatch.setenv("B200PT_OBJ_THREADS", str(threads))
        sc = pt.Scene(assets=assets, width=8, height=8, spp=1, depth=1)
        return [(e["name"], e["verts"].copy(), e["normals"].copy(), e["idx"].copy(), None if e["uvs"] is None else e["uvs"].copy())
                for e in sc.entities()]
    def same(a, b):
        assert [x[0] for x in a] == [x[0] for x in b] and len(a) > 0
        for x, y in zip(a, b):
            for u, v in zip(x[1:], y[1:]):
                assert (u is None) == (v is None)
                if u is not None:
                    assert u.shape == v.shape and (u.view(np.uint32) == v.view(np.uint32)).all()
    # (a) the reference's cornell box
    (tmp_path / "a").mkdir()
    for f in ("cornell-box.json", "cornell-box.obj.scene"):
        (tmp_path / "a" / f).write_bytes((REPO / "assets" / f).read_bytes())
    (tmp_path / "a" / "settings.json").write_text((REPO / "assets/settings.json").read_text().replace('"car"', '"cornell-box"'))
    serial = entities(tmp_path / "a", 1)
    assert sum(len(e[3]) for e in serial) == 17974
    for t in (2, 5, 16):
        same(serial, entities(tmp_path / "a", t))
    # (b) synthetic: 12 shapes, absolute and relative indices mixed, quads, CRLF, comments
    rng = np.random.default_rng(5)
    lines, nv, nn, nt = ["# synthetic"], 0, 0, 0
    for sidx in range(12):
        # trailing blanks only after `g` names: tinyobj keeps them in an `o` name (tests/test_ref_host.py)
        lines.append(("g " if sidx % 3 else "o ") + f"m{sidx}" + ("  " if sidx % 2 and sidx % 3 else ""))
        for _ in range(int(rng.integers(40, 400))):
            k = int(rng.integers(3, 6))
            for _ in range(k):
                x, y, z = rng.normal(size=3)
                lines.append(f"v {x:.6f} {y:.6f} {z:.6f}")
                lines.append(f"vn {x:.3f} {y:.3f} {z:.3f}")
                lines.append(f"vt {abs(x) % 1:.4f} {abs(y) % 1:.4f}")
            nv += k; nn += k; nt += k
            if rng.random() < 0.2:
                lines.append("# a comment between attributes and faces")
            for _ in range(int(rng.integers(1, 4))):
                corners = []
                for c in rng.permutation(k)[:int(rng.integers(3, min(k, 4) + 1))]:
                    if rng.random() < 0.5:   # relative: -1 is the attribute defined last
                        a = -(k - int(c))
                        corners.append(f"{a}/{a}/{a}" if rng.random() < 0.7 else f"{a}//{a}")
                    else:
                        a = nv - k + int(c) + 1
                        corners.append(f"{a}/{a}/{a}")
                lines.append("f " + " ".join(corners))
    text = "".join(l + ("\r\n" if i % 7 == 0 else "\n") for i, l in enumerate(lines))
    (tmp_path / "b").mkdir()
    (tmp_path / "b" / "t.obj.scene").write_text(text)
    (tmp_path / "b" / "settings.json").write_text((REPO / "assets/settings.json").read_text().replace('"car"', '"t"'))
    scene = json.loads((REPO / "assets/cube.json").read_text())
    proto = dict(scene["materials"][0], use_texture=False)
    scene["materials"] = [dict(proto, name=f"m{i}") for i in range(12)]
    (tmp_path / "b" / "t.json").write_text(json.dumps(scene))
    serial = entities(tmp_path / "b", 1)
    assert len(serial) == 12
    for t in (2, 3, 7, 16):
        same(serial, entities(tmp_path / "b", t))


def png_pixels(path):
    d = path.read_bytes()
    assert d[:8] == b"\x89PNG\r\n\x1a\n"
    pos, idat, w, h = 8, b"", 0, 0
    while pos < len(d):
        n, t = struct.unpack(">I4s", d[pos:pos + 8])
        body = d[pos + 8:pos + 8 + n]
        assert zlib.crc32(t + body) == struct.unpack(">I", d[pos + 8 + n:pos + 12 + n])[0]
        if t == b"IHDR":
            w, h = struct.unpack(">II", body[:8])
            assert body[8:13] == bytes([8, 6, 0, 0, 0])
        if t == b"IDAT":
            idat += body
        pos += 12 + n
    raw = np.frombuffer(zlib.decompress(idat), np.uint8).reshape(h, 1 + 4 * w)
    assert (raw[:, 0] == 0).all()
    return np.ascontiguousarray(raw[:, 1:]).view(np.uint32).reshape(h, w)


def test_pfm_writer(tmp_path):
    """Float frame next to the PNG (path_tracer --pfm): RGB, little endian, bottom row first."""
    H = pt.host_lib()
    rs = np.random.RandomState(1)
    img = rs.rand(5, 7, 4).astype(np.float32) * 30.0
    p = tmp_path / "x.pfm"
    assert H.b200pt_write_pfm(str(p).encode(), 7, 5, img.ctypes.data) == 0
    raw = p.read_bytes()
    head, w_h, scale, body = raw.split(b"\n", 3)
    assert head == b"PF" and w_h == b"7 5" and float(scale) < 0
    back = np.frombuffer(body, "<f4").reshape(5, 7, 3)[::-1]
    assert np.array_equal(back, img[..., :3])


def test_exr_writer(tmp_path):
    """path_tracer --exr: single-part scan-line OpenEXR, uncompressed, FLOAT channels in the format's B G R order,
    top row first; read back here with a reader written from the file layout (magic, attributes, offset table)."""
    import struct
    H = pt.host_lib()
    rs = np.random.RandomState(2)
    w, h = 9, 6
    img = (rs.rand(h, w, 4).astype(np.float32) - 0.2) * 1000.0
    p = tmp_path / "x.exr"
    assert H.b200pt_write_exr(str(p).encode(), w, h, img.ctypes.data) == 0
    raw = p.read_bytes()
    assert raw[:4] == bytes([0x76, 0x2F, 0x31, 0x01]) and struct.unpack("<I", raw[4:8])[0] == 2
    at, attrs = 8, {}
    while raw[at] != 0:
        e = raw.index(b"\0", at); name = raw[at:e].decode(); at = e + 1
        e = raw.index(b"\0", at); typ = raw[at:e].decode(); at = e + 1
        size = struct.unpack("<i", raw[at:at + 4])[0]; at += 4
        attrs[name] = (typ, raw[at:at + size]); at += size
    at += 1
    assert attrs["compression"] == ("compression", b"\0") and attrs["lineOrder"] == ("lineOrder", b"\0")
    assert struct.unpack("<4i", attrs["dataWindow"][1]) == (0, 0, w - 1, h - 1) == struct.unpack("<4i", attrs["displayWindow"][1])
    ch, names, c = attrs["channels"][1], [], 0
    while ch[c] != 0:
        e = ch.index(b"\0", c); names.append(ch[c:e].decode())
        assert struct.unpack("<iB3xii", ch[e + 1:e + 17]) == (2, 0, 1, 1)       # FLOAT, not linear-hinted, sampling 1 x 1
        c = e + 17
    assert names == ["B", "G", "R"]
    offsets = struct.unpack(f"<{h}Q", raw[at:at + 8 * h])
    back = np.empty((h, w, 3), np.float32)
    for y, off in enumerate(offsets):
        yy, size = struct.unpack("<ii", raw[off:off + 8])
        assert yy == y and size == w * 12
        planes = np.frombuffer(raw[off + 8:off + 8 + size], "<f4").reshape(3, w)
        back[y] = planes[::-1].T                                                # B G R planes -> R G B pixels
    assert offsets[-1] + 8 + w * 12 == len(raw)
    assert np.array_equal(back, img[..., :3])
    assert H.b200pt_write_exr(str(tmp_path / "no" / "dir.exr").encode(), w, h, img.ctypes.data) != 0


def test_png_writer_and_texture_reader_round_trip(tmp_path):
    H = pt.host_lib()
    rs = np.random.RandomState(0)
    img = rs.randint(0, 2 ** 32, (13, 17), dtype=np.uint64).astype(np.uint32) | np.uint32(0xFF000000)
    p = tmp_path / "x.png"
    assert H.b200pt_write_png(str(p).encode(), 17, 13, img.ctypes.data) == 0
    assert (png_pixels(p) == img).all()
    # same file as a base-colour texture: loaded, v-flipped (application.cpp:233-238)
    (tmp_path / "settings.json").write_text((REPO / "assets/settings.json").read_text().replace('"car"', '"cube"'))
    (tmp_path / "cube.json").write_text((REPO / "assets/cube.json").read_text().replace("cube.png", "x.png"))
    (tmp_path / "cube.obj.scene").write_text((REPO / "assets/cube.obj.scene").read_text())
    (tmp_path / "cube-textures").mkdir()
    (tmp_path / "cube-textures" / "x.png").write_bytes(p.read_bytes())
    sc = pt.Scene(assets=tmp_path, width=8, height=8, spp=1, depth=1)
    tex = sc.entities()[0]["tex"]
    assert (np.ascontiguousarray(tex).view(np.uint32)[..., 0] == img[::-1]).all()
    assert not sc.notes


def test_environment_hdr_goes_through_stb_ldr_route(tmp_path):
    """image_buffer.cpp:36-59 + stb_image.h:1864-1889: RGBE -> pow(x,1/2.2)*255+.5, v-flip."""
    w, h = 16, 4
    rgbe = np.zeros((h, w, 4), np.uint8)
    rgbe[..., 0] = 128
    rgbe[..., 1] = np.arange(w, dtype=np.uint8) * 8
    rgbe[..., 2] = 255
    rgbe[..., 3] = 128 + np.arange(h, dtype=np.uint8)[:, None] - 1
    with open(tmp_path / "environment.hdr", "wb") as f:
        f.write(b"#?RADIANCE\nFORMAT=32-bit_rle_rgbe\n\n-Y %d +X %d\n" % (h, w))
        for y in range(h):
            f.write(bytes([2, 2, 0, w]))
            for k in range(4):
                f.write(bytes([w]) + rgbe[y, :, k].tobytes())
    for n in ("settings.json", "cube.json", "cube.obj.scene"):
        (tmp_path / n).write_text((REPO / "assets" / n).read_text().replace('"car"', '"cube"'))
    sc = pt.Scene(assets=tmp_path, width=8, height=8, spp=1, depth=1, env="map")
    env = sc.environment["rgba"]
    assert env.shape == (h, w, 4)
    val = rgbe[..., :3].astype(np.float64) * np.exp2(rgbe[..., 3:4].astype(np.float64) - 136)
    ldr = np.clip((val.astype(np.float32).astype(np.float64) ** np.float64(np.float32(1 / 2.2))).astype(np.float32) * 255 + np.float32(0.5), 0, 255).astype(np.uint8)
    assert (env[..., :3] == ldr[::-1]).all() and (env[..., 3] == 255).all()


def test_material_sweep_host_side():
    sc = pt.Scene(scene="cornell-box", width=8, height=8, spp=1, depth=1)
    sc.set_material("box", "roughness", 0.75)
    sc.set_material("box", "base_color", [0.1, 0.2, 0.3])
    m = sc.materials()[sc.material_names().index("box")]
    assert m["roughness"] == np.float32(0.75) and m["base_color_g"] == np.float32(0.2)
    with pytest.raises(pt.B200ptError):
        sc.set_material("nope", "roughness", 1.0)


def test_procedural_stand_ins_are_deterministic_and_outward():
    a = pt.Scene(scene="sphere", width=8, height=8, spp=1, depth=1, env="auto")
    b = pt.Scene(scene="sphere", width=8, height=8, spp=1, depth=1, env="auto")
    ea, eb = a.entities()[0], b.entities()[0]
    assert ea["idx"].shape[0] == 16384 and (ea["verts"] == eb["verts"]).all()
    out = ea["verts"] - np.array([0, 1, 0], np.float32)
    assert (np.sum(out * ea["normals"], 1) > 0.9).mean() > 0.99      # normals point outward
    d = pt.Scene(scene="dragon", width=8, height=8, spp=1, depth=1)
    assert [e["name"] for e in d.entities()] == ["dragon", "ground", "areaLight"]
    assert d.n_triangles == 871406 and "procedural" in d.notes[0]
    g = d.entities()[1]
    assert (g["normals"][:, 1] > 0.99).all()                          # ground faces up
    assert (d.entities()[2]["normals"][:, 1] < -0.99).all()           # light faces down
