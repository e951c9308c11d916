"""binary32 vs binary64 deterministic trig (PT_TRIG_F32): how far is the GPU frame from the glibc-math oracle, per pixel,
on a crop of the headline frame?   python tools/trig_probe.py <libb200pt.so>"""
import json, os, sys, time
from pathlib import Path
R = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(R)); sys.path.insert(0, str(R / "tests"))
os.environ.setdefault("B200PT_QUIET", "1")
import numpy as np
import owl_path_tracer_b200 as pt
pt.use_library(sys.argv[1])
from checkers import Oracle, rel_mse
for name, spp in (("dragon", 1024), ("dragon", 256), ("car", 1024)):
    sc = pt.Scene(scene=name, width=1920, height=1080, spp=spp, depth=8, env="color", env_intensity=0.0)
    ctx = pt.Context(); ctx.bind(sc); ctx.render()
    g = ctx.framebuffer_f32()[..., :3]; ctx.close()
    x0, y0, cw, ch = 896, 476, 128, 128
    r = Oracle(det=False).load_scene(sc).render(window=(x0, y0, x0 + cw, y0 + ch))
    want = r["sum"][y0:y0 + ch, x0:x0 + cw, :3] * (np.float32(1.0) / np.float32(spp))
    got = g[::-1][y0:y0 + ch, x0:x0 + cw]
    print(json.dumps(dict(lib=Path(sys.argv[1]).name, scene=name, spp=spp, rel_mse=rel_mse(got, want),
                          pixels_differing=float(np.mean(np.any(got != want, axis=-1))))), flush=True)
