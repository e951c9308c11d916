"""A/B of variant builds of libb200pt.so (tools/build_variants.sh): every variant runs in its own process, bound explicitly
with owl_path_tracer_b200.use_library (never through the environment), on the same frames.

    python tools/ab.py [--spp 32] [--scenes dragon,car] [--timing] [--check] base=owl_path_tracer_b200/lib/libb200pt.so name=path ...

Prints one JSON line per (variant, scene): frame ms (product configuration: two concurrent sub-frames, no per-kernel
events; best of 3), Mrays/s, and with --timing the per-kernel times of a one-stream pass with events.  --check also
compares a 640x360 / 4 spp dragon frame (a size at which walks get parked) bit for bit with the oracle."""
import json
import os
import subprocess
import sys
from pathlib import Path

R = Path(__file__).resolve().parent.parent


def child(lib, scenes, spp, timing, check):
    os.environ.setdefault("B200PT_QUIET", "1")
    sys.path.insert(0, str(R)); sys.path.insert(0, str(R / "tests"))
    import numpy as np
    import owl_path_tracer_b200 as pt
    pt.use_library(lib)
    out = []
    if check:
        from checkers import Oracle
        sc = pt.Scene(scene="dragon", width=640, height=360, spp=4, depth=8, env="color", env_intensity=0.0)
        ctx = pt.Context(); ctx.bind(sc); ctx.render()
        ro = Oracle(det=True).load_scene(sc).render()
        same = (np.ascontiguousarray(ctx.framebuffer_f32()[..., :3]).view(np.uint32) == np.ascontiguousarray(ro["mean"][..., :3]).view(np.uint32)).all()
        out.append(dict(check="dragon 640x360 4spp", bit_equal=bool(same), rays_equal=ctx.stats()["rays"] == ro["rays"]))
        ctx.close()
    for scene in scenes:
        sc = pt.Scene(scene=scene, width=1920, height=1080, spp=spp, depth=8, env="color", env_intensity=0.0)
        ctx = pt.Context(); ctx.bind(sc)
        ms = []
        for _ in range(4):
            ctx.render(); st = ctx.stats(); ms.append(st["render_ms"])
        row = dict(scene=scene, spp=spp, frame_ms=min(ms[1:]), mrays=st["rays"] / min(ms[1:]) * 1e-3, rays=st["rays"], iters=st["iterations"], build_ms=st.get("build_ms"), nodes=st.get("bvh_nodes"))
        ctx.close()
        if timing:
            c = pt.Context(flags=pt.FLAG_TIMING | pt.FLAG_COUNTERS); c.bind(sc); c.render(); c.render(); t = c.stats()
            row.update(extend_ms=t["extend_ms"], shade_ms=t["shade_ms"], serial_ms=t["render_ms"], nodes_per_ray=t["nodes_visited"] / t["rays"],
                       tris_per_ray=t["tris_tested"] / t["rays"])
            c.close()
        out.append(row)
    print(json.dumps(out))


def main():
    args = sys.argv[1:]
    if args and args[0] == "--child":
        return child(args[1], args[2].split(","), int(args[3]), args[4] == "1", args[5] == "1")
    spp, scenes, timing, check, variants = 32, "dragon,car", False, False, []
    i = 0
    while i < len(args):
        a = args[i]
        if a == "--spp": spp = int(args[i + 1]); i += 2
        elif a == "--scenes": scenes = args[i + 1]; i += 2
        elif a == "--timing": timing = True; i += 1
        elif a == "--check": check = True; i += 1
        else: variants.append(a.split("=", 1)); i += 1
    for name, lib in variants:
        r = subprocess.run([sys.executable, __file__, "--child", str(Path(lib).resolve()), scenes, str(spp), "1" if timing else "0", "1" if check else "0"],
                           capture_output=True, text=True, timeout=900)
        if r.returncode != 0:
            print(json.dumps(dict(variant=name, error=r.stderr[-400:])), flush=True)
            continue
        for row in json.loads(r.stdout.strip().splitlines()[-1]):
            print(json.dumps(dict(variant=name, **row)), flush=True)


if __name__ == "__main__":
    main()
