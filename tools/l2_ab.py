"""L2 experiments: persisting window on the triangle blocks / nodes (B200PT_L2_PERSIST) in one process."""
import json, os, sys
from pathlib import Path
R = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(R)); sys.path.insert(0, str(R / "tests"))
os.environ.setdefault("B200PT_QUIET", "1")
import numpy as np
import owl_path_tracer_b200 as pt
for scene in ("dragon", "car"):
    sc = pt.Scene(scene=scene, width=1920, height=1080, spp=128, depth=8, env="color", env_intensity=0.0)
    ctx = pt.Context(); ctx.bind(sc)
    base = None
    for mode in (0, 1, 2, 0, 1):
        os.environ["B200PT_L2_PERSIST"] = str(mode)
        ctx.render(); ctx.render()
        st = ctx.stats(); f = ctx.framebuffer_f32()
        if base is None: base = f.copy()
        print(json.dumps(dict(scene=scene, l2_persist=mode, render_ms=st["render_ms"], mrays=st["rays"] / st["render_ms"] * 1e-3,
                              same_frame=bool((f.view(np.uint32) == base.view(np.uint32)).all()))), flush=True)
    ctx.close()
