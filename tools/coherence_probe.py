#!/usr/bin/env python
"""How much does ray ORDER matter to the walker?  Second-bounce-like rays of the 1080p dragon frame
(origins = primary hit points in pixel order, directions = random hemisphere around the facet normal)
are traced through b200pt_trace in different orders; stats.render_ms is the walk kernel alone."""
import json, os, sys
from pathlib import Path
R = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(R)); sys.path.insert(0, str(R / "tests"))
os.environ.setdefault("B200PT_QUIET", "1")
import numpy as np
import owl_path_tracer_b200 as pt


def main():
    scene = sys.argv[1] if len(sys.argv) > 1 else "dragon"
    W, H = 1920, 1080
    sc = pt.Scene(scene=scene, width=W, height=H, spp=1, depth=8, env="color", env_intensity=0.0)
    ctx = pt.Context(); ctx.bind(sc)
    org, llc, hor, ver = sc.camera
    ys, xs = np.mgrid[0:H, 0:W]
    su, sv = ((xs + 0.5) / W).astype(np.float32)[..., None], ((ys + 0.5) / H).astype(np.float32)[..., None]
    D0 = (llc + su * hor + sv * ver - org).reshape(-1, 3).astype(np.float32)
    D0 /= np.linalg.norm(D0, axis=1, keepdims=True)
    O0 = np.broadcast_to(org, D0.shape).astype(np.float32).copy()
    h0 = ctx.trace(O0, D0); t_primary = ctx.stats()["render_ms"]
    hit = h0["mesh"] >= 0
    P = O0 + D0 * h0["t"][:, None]
    ents = sc.entities()
    N = np.zeros_like(P)
    for m, en in enumerate(ents):
        sel = np.nonzero(h0["mesh"] == m)[0]
        if sel.size == 0: continue
        tri = en["verts"][en["idx"][h0["prim"][sel]]]
        n = np.cross(tri[:, 1] - tri[:, 0], tri[:, 2] - tri[:, 0])
        n /= np.maximum(np.linalg.norm(n, axis=1, keepdims=True), 1e-30)
        n[(n * D0[sel]).sum(1) > 0] *= -1       # face the camera
        N[sel] = n
    rs = np.random.RandomState(1)
    V = rs.normal(size=P.shape).astype(np.float32); V /= np.linalg.norm(V, axis=1, keepdims=True)
    V[(V * N).sum(1) < 0] *= -1
    O1, D1 = (P + 1e-3 * N)[hit].astype(np.float32), V[hit].astype(np.float32)
    n = O1.shape[0]
    out = dict(scene=scene, primary=dict(rays=int(D0.shape[0]), ms=t_primary, mrays=D0.shape[0] / t_primary * 1e-3), bounce_rays=n)

    def run(tag, order):
        o, d = np.ascontiguousarray(O1[order]), np.ascontiguousarray(D1[order])
        ctx.trace(o, d); ctx.trace(o, d)
        ms = ctx.stats()["render_ms"]
        out[tag] = dict(ms=ms, mrays=n / ms * 1e-3)
        print(tag, out[tag], flush=True)

    idx = np.arange(n)
    octant = (D1[:, 0] < 0) * 4 + (D1[:, 1] < 0) * 2 + (D1[:, 2] < 0) * 1
    lo, hi = O1.min(0), O1.max(0)
    q = np.clip(((O1 - lo) / np.maximum(hi - lo, 1e-9) * 1023).astype(np.uint64), 0, 1023)
    def spread(v):
        v = (v | (v << 16)) & 0x030000FF; v = (v | (v << 8)) & 0x0300F00F
        v = (v | (v << 4)) & 0x030C30C3; v = (v | (v << 2)) & 0x09249249
        return v
    morton = (spread(q[:, 0]) << 2) | (spread(q[:, 1]) << 1) | spread(q[:, 2])
    run("pixel_order", idx)
    run("shuffled", rs.permutation(n))
    run("by_octant", np.argsort(octant, kind="stable"))
    run("by_octant_then_morton", np.lexsort((morton, octant)))
    run("by_morton_then_octant", np.lexsort((octant, morton >> 15)))
    # tiles: 8x4-pixel tiles are already contiguous per warp? (pixel order = rows) -> tile order
    px = np.nonzero(hit)[0]; x, y = px % W, px // W
    run("tiles_8x4", np.lexsort((x % 8, y % 4, x // 8, y // 4)))
    run("tiles_8x4_by_octant", np.lexsort((octant, x // 8, y // 4)))
    # throughput against launch size: random subsets of the same ray population
    for frac in (1.0, 0.5, 0.25, 0.125, 0.0625, 0.03125):
        k = int(n * frac)
        sub = rs.permutation(n)[:k]
        o, d = np.ascontiguousarray(O1[sub]), np.ascontiguousarray(D1[sub])
        ctx.trace(o, d); ctx.trace(o, d)
        ms = ctx.stats()["render_ms"]
        out[f"subset_{k}"] = dict(ms=ms, mrays=k / ms * 1e-3)
        print("subset", k, out[f"subset_{k}"], flush=True)
    print(json.dumps(out))
    (R / "gpurun_out").mkdir(exist_ok=True)
    (R / "gpurun_out" / f"coherence_{scene}.json").write_text(json.dumps(out, indent=1))
    ctx.close()


if __name__ == "__main__":
    main()
