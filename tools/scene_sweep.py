#!/usr/bin/env python
"""Throughput of every BASELINE.json config on one B200 + image agreement with the CPU oracle.

For each bundled scene (cornell-box, sphere, cube, mitsuba, car, dragon) at the config's size:
  * Mrays/s and spp/s of b200pt_render at the config's resolution / spp (device time, CUDA events);
  * relMSE of the GPU frame against the glibc-math oracle (independent rounding -> different paths,
    same estimator) at MATCHED spp on a reduced frame (the oracle is scalar CPU code), compared after
    8x8 box filtering, plus the bit-equality check against the deterministic-math oracle.
Writes one JSON line per scene to stdout (and to gpurun_out/scene_sweep.jsonl).
"""
import json
import os
import sys
import time
from pathlib import Path

R = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(R))
sys.path.insert(0, str(R / "tests"))
os.environ.setdefault("B200PT_QUIET", "1")
import numpy as np  # noqa: E402
import owl_path_tracer_b200 as pt  # noqa: E402
from checkers import Oracle, rel_mse  # noqa: E402

CONFIGS = [  # (config id, scene, W, H, spp, depth, env kwargs)
    ("configs[0]", "cornell-box", 512, 512, 64, 8, dict(env="color", env_intensity=0.0)),
    ("configs[1]", "sphere", 1024, 1024, 256, 8, dict(env="map", env_intensity=1.0)),
    ("configs[1]", "cube", 1024, 1024, 256, 8, dict(env="map", env_intensity=1.0)),
    ("configs[2]", "mitsuba", 1920, 1080, 512, 8, dict(env="auto", env_intensity=1.0)),
    ("configs[3]", "car", 1920, 1080, 1024, 8, dict(env="color", env_intensity=0.0)),
    ("configs[4]", "dragon", 1920, 1080, 1024, 8, dict(env="color", env_intensity=0.0)),
]


def bits(a):
    return np.ascontiguousarray(a).view(np.uint32)


def pool(a, k=8):
    h, w = a.shape[0] // k * k, a.shape[1] // k * k
    return a[:h, :w].reshape(h // k, k, w // k, k, 3).mean((1, 3))


def main():
    quick = "--quick" in sys.argv
    out = []
    for cfg, scene, W, H, spp, depth, kw in CONFIGS:
        full_spp = spp if not quick else min(spp, 32)
        sc = pt.Scene(scene=scene, width=W, height=H, spp=full_spp, depth=depth, **kw)
        ctx = pt.Context(flags=pt.FLAG_TIMING)
        ctx.bind(sc)
        ctx.render()                      # warm-up (also touches every page)
        ctx.render()
        st = ctx.stats()
        ctx.close()
        # agreement at matched spp on a reduced frame (oracle cost ~ 6 Mrays/s on 16 cores)
        w2, h2 = (256, 256) if W == H else (320, 180)
        spp2 = min(spp, 64)
        sc2 = pt.Scene(scene=scene, width=w2, height=h2, spp=spp2, depth=depth, **kw)
        c2 = pt.Context()
        c2.bind(sc2)
        c2.render()
        g = c2.framebuffer_f32()[..., :3].copy()
        c2.close()
        t0 = time.time()
        o = Oracle(det=False).load_scene(sc2).render()["mean"][..., :3]
        t_or = time.time() - t0
        od = Oracle(det=True).load_scene(sc2).render()["mean"][..., :3]
        line = dict(config=cfg, scene=scene, size=[W, H], spp=full_spp, depth=depth, triangles=sc.n_triangles,
                    mrays_per_s=st["rays"] / st["render_ms"] * 1e-3, spp_per_s=full_spp / (st["render_ms"] * 1e-3),
                    render_ms=st["render_ms"], extend_ms=st["extend_ms"], shade_ms=st["shade_ms"], rays=st["rays"],
                    rays_per_path=st["rays"] / st["paths"], bvh_nodes=st["bvh_nodes"], build_ms=st["build_ms"],
                    check=dict(size=[w2, h2], spp=spp2, relmse_vs_glibc_oracle_8x8=rel_mse(pool(g), pool(o)),
                               relmse_vs_glibc_oracle_pixel=rel_mse(g, o), mean_gpu=float(g.mean()), mean_oracle=float(o.mean()),
                               bit_equal_to_det_oracle=bool((bits(g) == bits(od)).all()), oracle_seconds=t_or),
                    notes=sc.notes)
        out.append(line)
        print(json.dumps(line), flush=True)
    (R / "gpurun_out").mkdir(exist_ok=True)
    (R / "gpurun_out" / "scene_sweep.jsonl").write_text("".join(json.dumps(x) + "\n" for x in out))


if __name__ == "__main__":
    main()
