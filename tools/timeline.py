"""Per-warp timeline of single k_extend launches.  Needs a development build (tools/build_variants.sh timeline
"-DPT_TIMELINE=1 -DB200PT_DEV=1"), passed as the first argument: python tools/timeline.py <path/to/libb200pt_timeline.so>:
when does each persistent warp start, see the queue drained, and exit; how many rays and loop steps it ran."""
import json, os, sys
from pathlib import Path
R = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(R)); sys.path.insert(0, str(R / "tests"))
os.environ.setdefault("B200PT_QUIET", "1")
import numpy as np
import owl_path_tracer_b200 as pt
pt.use_library(sys.argv[1])
out = R / "gpurun_out"; out.mkdir(exist_ok=True)
sc = pt.Scene(scene="dragon", width=1920, height=1080, spp=32, depth=8, env="color", env_intensity=0.0)
ctx = pt.Context(flags=pt.FLAG_TIMING); ctx.bind(sc); ctx.render()
os.environ["B200PT_ITER_LOG"] = str(out / "timeline_iters.txt")
ctx.render()
os.environ.pop("B200PT_ITER_LOG")
iters = np.loadtxt(out / "timeline_iters.txt")
for it in (10, 45, 75, 100, 130, 170, 200):
    if it >= len(iters): continue
    f = out / f"timeline_{it}.txt"
    os.environ["B200PT_TIMELINE_ITER"] = str(it); os.environ["B200PT_TIMELINE_FILE"] = str(f)
    ctx.render()
    a = np.loadtxt(f)
    if a.ndim != 2 or not len(a):
        print(json.dumps(dict(iter=it, error="no records"))); continue
    t0 = a[:, 1].min()
    start, drained, exit_, rays, steps = (a[:, 1] - t0) / 1e3, (a[:, 2] - t0) / 1e3, (a[:, 3] - t0) / 1e3, a[:, 4], a[:, 5]
    node_lanes, tri_lanes, node_steps = a[:, 6].sum(), a[:, 7].sum(), a[:, 8].sum()
    tri_steps = steps.sum() - node_steps
    q = lambda x, p: float(np.percentile(x, p))
    print(json.dumps(dict(iter=it, queue=int(iters[it][0]), parked=int(iters[it][1]), extend_ms=float(iters[it][2]), warps=len(a), rays=int(rays.sum()),
        start_us=dict(p50=q(start, 50), p99=q(start, 99), max=float(start.max())),
        drained_us=dict(min=float(drained.min()), p50=q(drained, 50), max=float(drained.max())),
        exit_us=dict(p10=q(exit_, 10), p50=q(exit_, 50), p90=q(exit_, 90), p99=q(exit_, 99), max=float(exit_.max())),
        node_steps_per_ray=node_lanes / rays.sum(), tri_steps_per_ray=tri_lanes / rays.sum(),
        lanes_per_node_step=node_lanes / max(node_steps, 1), lanes_per_tri_step=tri_lanes / max(tri_steps, 1),
        steps=dict(p50=q(steps, 50), max=float(steps.max())), rays_per_warp=dict(p50=q(rays, 50), max=float(rays.max())))), flush=True)
ctx.close()
