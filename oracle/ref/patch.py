#!/usr/bin/env python3
"""oracle/ref/patch.py — copies the reference sources that make up the render path into a build
directory (oracle/_ref/build, git-ignored) and applies the TWO mechanical edits g++ needs.

TEST INFRASTRUCTURE.  No reference text is ever written into the tracked tree.

The edit (random.hpp:58-75): the two explicit specialisations `template<> ... rng()` are written
inside the class body, which MSVC/nvcc accept and g++ rejects (GCC bug 85282).  They are moved,
token for token, behind the class as `random::rng<float>` / `random::rng<vec2>`.  Nothing else in any
file is changed; the name clash of `struct random` with POSIX random(3), the CUDA qualifiers and the
unqualified ::max are handled from outside by shim/ref_prelude.h.

The second edit (utils/parser.cpp:26,69-70,85,88): `auto config{load_json(...)}` / `auto const cam{config["camera"]}` /
`auto test{config["test"]}` brace-initialise a nlohmann::json from a json.  MSVC (the reference's compiler) copies; g++
prefers the initializer_list constructor and builds a one-element ARRAY around the value (a documented nlohmann/GCC
difference), after which every `config["..."]` throws type_error.305.  The braces become `=`: the MSVC meaning.

Also extracts make_8bit / make_rgba (extern/owl/owl/include/owl/owl_device.h:79-91) into
gen/owl_make_rgba.inc so the shim's owl_device.h can include the real quantisation.
"""
import re
import shutil
import sys
from pathlib import Path

REF = Path(sys.argv[1])
OUT = Path(sys.argv[2])
SRC = REF / "path_tracer" / "src"

FILES = [
    "types.hpp", "macros.hpp", "random.hpp", "math.hpp", "sample_methods.hpp", "camera.hpp", "camera.cpp",
    "device/device.hpp", "device/device_global.hpp", "device/device.cu", "device/lights.hpp",
    "device/disney/disney.cuh", "device/disney/disney_helper.cuh", "device/disney/disney_diffuse.cuh",
    "device/disney/disney_sheen.cuh", "device/disney/disney_clearcoat.cuh", "device/disney/disney_specular.cuh",
    "utils/mesh_loader.hpp", "utils/mesh_loader.cpp", "utils/image_buffer.hpp", "utils/image_buffer.cpp",
    "utils/parser.hpp", "utils/parser.cpp",
]


def move_specialisations(text: str) -> str:
    """random.hpp: cut `template<> inline __both__ T rng() {...}` blocks out of the class, paste them after it."""
    pat = re.compile(r"\n    template<>\n    inline __both__ (\w+) rng\(\)\n    \{.*?\n    \}\n", re.S)
    blocks = pat.findall(text)
    moved = []
    for m in pat.finditer(text):
        body = m.group(0)
        ty = m.group(1)
        moved.append(body.replace(f"inline __both__ {ty} rng()", f"inline __both__ {ty} random::rng<{ty}>()"))
    assert len(moved) == 2 and blocks == ["float", "vec2"], "random.hpp no longer has the two in-class specialisations"
    text = pat.sub("\n", text)
    end = text.index("\n};", text.index("struct random")) + 3     # the class's own closing brace (column 0)
    return text[:end] + "\n" + "".join(moved) + text[end:]


def main():
    if OUT.exists():
        shutil.rmtree(OUT)
    for f in FILES:
        dst = OUT / "src" / f
        dst.parent.mkdir(parents=True, exist_ok=True)
        text = (SRC / f).read_text(encoding="utf-8-sig")
        if f == "random.hpp":
            text = move_specialisations(text)
        if f == "utils/parser.cpp":
            text, n = re.subn(r'auto( const)? (\w+)\{(load_json\(config_path\)|config\["\w+"\])\};', r'auto\1 \2 = \3;', text)
            assert n == 5, f"parser.cpp: expected 5 brace-initialised json copies, found {n}"
        dst.write_text(text)
    gen = OUT / "gen"
    gen.mkdir(parents=True, exist_ok=True)
    lines = (REF / "extern/owl/owl/include/owl/owl_device.h").read_text().splitlines()
    start = next(i for i, l in enumerate(lines) if "uint32_t make_8bit(const float f)" in l)
    stop = next(i for i, l in enumerate(lines) if "uint32_t make_rgba(const vec4f color)" in l)
    (gen / "owl_make_rgba.inc").write_text("\n".join(lines[start:stop]) + "\n")


if __name__ == "__main__":
    main()
