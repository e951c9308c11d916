"""CPU-side BVH quality probe: builds the tree with the emulated device builder (same per-element
code as bvh_build.cu) and counts node fetches / triangle tests per ray over a small render."""
import sys, time, json
from pathlib import Path
R = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(R)); sys.path.insert(0, str(R / "tests"))
import os
os.environ.setdefault("B200PT_QUIET", "1")
import owl_path_tracer_b200 as pt
from checkers import Emu

def probe(scene, w=320, h=180, spp=2, **kw):
    sc = pt.Scene(scene=scene, width=w, height=h, spp=spp, depth=8, **kw)
    e = Emu().load_scene(sc)
    t = time.time(); n = e.build(); tb = time.time() - t   # load_scene already built once; timed rebuild
    r = e.render()
    print(json.dumps(dict(scene=scene, tris=sc.n_triangles, nodes=n, levels=e.levels, build_s=round(tb, 2), rays=r["rays"],
                          nodes_per_ray=round(r["nodes_visited"] / r["rays"], 3), tris_per_ray=round(r["tris_tested"] / r["rays"], 3),
                          cost=round((4.4 * r["nodes_visited"] + r["tris_tested"]) / r["rays"], 2))))

if __name__ == "__main__":
    scenes = sys.argv[1:] or ["dragon", "car", "cornell-box"]
    for s in scenes:
        probe(s, env="color", env_intensity=0.0)
