"""GPU frames against the glibc-math oracle (independent of the product's deterministic functions), per pixel, unpooled.
Prints relMSE per case: the numbers behind the gates of tests/test_independent_parity.py."""
import json, os, sys, time
from pathlib import Path
R = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(R)); sys.path.insert(0, str(R / "tests"))
os.environ.setdefault("B200PT_QUIET", "1")
import numpy as np
import owl_path_tracer_b200 as pt
from checkers import Oracle, rel_mse


def case(name, w, h, spp, crop=None, sweep=None, **kw):
    sc = pt.Scene(scene=name, width=w, height=h, spp=spp, depth=8, **kw)
    if sweep:
        sc.set_material(*sweep)
    ctx = pt.Context(); ctx.bind(sc); ctx.render()
    g = ctx.framebuffer_f32()[..., :3]
    rays = ctx.stats()["rays"]
    ctx.close()
    o = Oracle(det=False).load_scene(sc)
    t0 = time.time()
    if crop:
        x0, y0, cw, ch = crop
        r = o.render(window=(x0, y0, x0 + cw, y0 + ch))
        want = r["sum"][y0:y0 + ch, x0:x0 + cw, :3] * (np.float32(1.0) / np.float32(spp))
        got = g[::-1][y0:y0 + ch, x0:x0 + cw]
    else:
        r = o.render()
        want, got = r["mean"][..., :3], g
    differ = float(np.mean(np.any(got != want, axis=-1)))
    print(json.dumps(dict(scene=name, size=[w, h], spp=spp, crop=crop, sweep=sweep, rel_mse=rel_mse(got, want), pixels_differing=differ,
                          mean_gpu=float(got.mean()), mean_oracle=float(want.mean()), oracle_s=round(time.time() - t0, 1))), flush=True)


case("car", 480, 270, 64, env="color", env_intensity=0.0)
case("sphere", 256, 256, 64, env="map", env_intensity=1.0)
case("cube", 256, 256, 64, env="auto", env_intensity=1.0)
for sw in (("outside", "metallic", 1.0), ("outside", "roughness", 0.1), ("outside", "clearcoat", 1.0), ("outside", "specular_transmission", 1.0),
           ("outside", "sheen", 1.0), ("outside", "anisotropic", 0.8)):
    case("mitsuba", 320, 180, 32, sweep=sw, env="auto", env_intensity=1.0)
case("dragon", 1920, 1080, 1024, crop=(832, 412, 256, 256), env="color", env_intensity=0.0)
case("car", 1920, 1080, 1024, crop=(832, 412, 256, 256), env="color", env_intensity=0.0)
