"""Fused tail kernel (k_fused): frame time against the queue length below which it takes over (B200PT_FUSE_MIN)
and its segments per launch (B200PT_FUSE_SEGS); parity of one switched frame against the oracle first."""
import json, os, sys
from pathlib import Path
R = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(R)); sys.path.insert(0, str(R / "tests"))
os.environ.setdefault("B200PT_QUIET", "1")
import numpy as np
import owl_path_tracer_b200 as pt
from checkers import Oracle

def bits(a): return np.ascontiguousarray(a).view(np.uint32)

# parity: a frame large enough to park walks, switched to the fused kernel half way
os.environ["B200PT_FUSE_MIN"] = "150000"; os.environ["B200PT_FUSE_SEGS"] = "5"
sc = pt.Scene(scene="dragon", width=640, height=360, spp=4, depth=8, env="color", env_intensity=0.0)
ctx = pt.Context(); ctx.bind(sc); ctx.render()
ro = Oracle(det=True).load_scene(sc).render()
print(json.dumps(dict(parity_dragon_640x360=dict(frame_bit_mismatch=int((bits(ctx.framebuffer_f32()[..., :3]) != bits(ro["mean"][..., :3])).sum()),
      rays_gpu=ctx.stats()["rays"], rays_oracle=ro["rays"], launches=ctx.stats()["kernel_launches"]))), flush=True)
ctx.close()

cases = [("dragon", 1920, 1080, 128), ("car", 1920, 1080, 128), ("cornell-box", 512, 512, 64)]
if len(sys.argv) > 1 and sys.argv[1] == "long":
    cases = [("dragon", 1920, 1080, 1024)]
for scene, w, h, spp in cases:
    sc = pt.Scene(scene=scene, width=w, height=h, spp=spp, depth=8, env="color", env_intensity=0.0)
    ctx = pt.Context(); ctx.bind(sc)
    base = None
    for fmin, segs in [(0, 8), (20000, 8), (50000, 8), (100000, 4), (100000, 8), (100000, 16), (200000, 8), (400000, 8)]:
        os.environ["B200PT_FUSE_MIN"] = str(fmin); os.environ["B200PT_FUSE_SEGS"] = str(segs)
        ctx.render(); ctx.render()
        st = ctx.stats()
        f = ctx.framebuffer_f32()
        if base is None: base = f.copy()
        print(json.dumps(dict(scene=scene, spp=spp, fuse_min=fmin, segs=segs, render_ms=st["render_ms"], mrays=st["rays"] / st["render_ms"] * 1e-3,
                              rays=st["rays"], launches=st["kernel_launches"], same_frame=bool((bits(f) == bits(base)).all()))), flush=True)
    ctx.close()
