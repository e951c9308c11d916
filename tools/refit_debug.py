import json, os, sys
from pathlib import Path
R = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(R)); sys.path.insert(0, str(R / "tests"))
os.environ.setdefault("B200PT_QUIET", "1")
import numpy as np
import owl_path_tracer_b200 as pt
from checkers import Oracle
from test_refit import moved, stress_rays, same_hits, checker_with

sc = pt.Scene(scene="dragon", width=320, height=180, spp=1, depth=8, env="color", env_intensity=0.0)
ents = sc.entities()
which = max(range(len(ents)), key=lambda i: len(ents[i]["idx"]))
print("meshes", [(i, len(e["idx"])) for i, e in enumerate(ents)], "which", which)
ctx = pt.Context(); ctx.bind(sc)
org0, d0 = stress_rays(ents, which, 200000, 3)
o0 = checker_with(Oracle, sc, ents, det=True)
print("unmoved: ctx vs oracle mismatches", int((~same_hits(ctx.trace(org0, d0), o0.trace(org0, d0))).sum()))
ctx.refit_accel()
print("unmoved after refit: mismatches", int((~same_hits(ctx.trace(org0, d0), o0.trace(org0, d0))).sum()))
new = moved(ents, which, scale=1.02, shift=(0.01, 0.03, 0.0), wobble=0.002)
ctx.update_mesh(which, new[which]["verts"]); ctx.refit_accel()
fresh = pt.Context()
for e in new:
    fresh.add_mesh(e["verts"], e["normals"], e["idx"], e["material"], e["uvs"], e["tex"])
fresh.build_accel(); fresh.set_materials(sc.materials()); fresh.set_camera(*sc.camera); fresh.set_render(sc.width, sc.height, sc.spp, sc.depth)
org, d = stress_rays(new, which, 200000, 3)
o = checker_with(Oracle, sc, new, det=True)
a, b, r = ctx.trace(org, d), fresh.trace(org, d), o.trace(org, d)
ma, mb = ~same_hits(a, r), ~same_hits(b, r)
print("moved: refit vs oracle", int(ma.sum()), " fresh vs oracle", int(mb.sum()), " refit vs fresh", int((~same_hits(a, b)).sum()))
for i in np.nonzero(ma | mb)[:1][0][:8]:
    print(i, "refit", a[i], "fresh", b[i], "oracle", r[i], "brute", o.trace(org[i:i+1], d[i:i+1], brute=True)[0])
