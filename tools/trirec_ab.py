"""Per-triangle shading records (SceneView::tri_rec) against the idx4 + vertex gathers (B200PT_NO_TRIREC=1), one process."""
import json, os, sys
from pathlib import Path
R = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(R)); sys.path.insert(0, str(R / "tests"))
os.environ.setdefault("B200PT_QUIET", "1")
import numpy as np
import owl_path_tracer_b200 as pt
for scene, w, h, spp, env in (("dragon", 1920, 1080, 128, "color"), ("car", 1920, 1080, 128, "color"), ("cornell-box", 512, 512, 64, "color"), ("mitsuba", 1920, 1080, 64, "auto")):
    sc = pt.Scene(scene=scene, width=w, height=h, spp=spp, depth=8, env=env, env_intensity=1.0 if env == "auto" else 0.0)
    for flags, tag in ((0, "product"), (pt.FLAG_TIMING, "timing")):
        ctx = pt.Context(flags=flags); ctx.bind(sc)
        base = None
        for off in (True, False, True, False):
            if off: os.environ["B200PT_NO_TRIREC"] = "1"
            else: os.environ.pop("B200PT_NO_TRIREC", None)
            ctx.render(); ctx.render()
            st = ctx.stats(); f = ctx.framebuffer_f32()
            if base is None: base = f.copy()
            print(json.dumps(dict(scene=scene, mode=tag, tri_records=not off, render_ms=st["render_ms"], extend_ms=st["extend_ms"], shade_ms=st["shade_ms"],
                                  mrays=st["rays"] / st["render_ms"] * 1e-3, same_frame=bool((f.view(np.uint32) == base.view(np.uint32)).all()))), flush=True)
        ctx.close()
os.environ.pop("B200PT_NO_TRIREC", None)
