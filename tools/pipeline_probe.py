"""Does rendering frame k+1 on a second context while frame k drains pay?  Two contexts bound to the same scene, one host
thread each (ctypes releases the GIL during b200pt_render), the second started `stagger` of a frame time after the first;
every thread renders `frames` frames back to back.  Prints frames/s against one context alone."""
import json, os, sys, threading, time
from pathlib import Path
R = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(R))
os.environ.setdefault("B200PT_QUIET", "1")
import owl_path_tracer_b200 as pt

spp = int(sys.argv[1]) if len(sys.argv) > 1 else 128
scene = sys.argv[2] if len(sys.argv) > 2 else "dragon"
frames = 4
sc = pt.Scene(scene=scene, width=1920, height=1080, spp=spp, depth=8, env="color", env_intensity=0.0)
ctxs = [pt.Context() for _ in range(3)]
for c in ctxs:
    c.bind(sc); c.render()
t0 = time.perf_counter()
for _ in range(frames):
    ctxs[0].render()
single = (time.perf_counter() - t0) / frames
rays = ctxs[0].stats()["rays"]
print(json.dumps(dict(scene=scene, spp=spp, mode="one context", ms_per_frame=single * 1e3, mrays=rays / single * 1e-6)), flush=True)
for n_ctx, stagger in ((2, 0.0), (2, 0.5), (3, 1 / 3)):
    def work(c, delay):
        time.sleep(delay)
        for _ in range(frames):
            c.render()
    th = [threading.Thread(target=work, args=(ctxs[i], i * stagger * single)) for i in range(n_ctx)]
    t0 = time.perf_counter()
    for t in th: t.start()
    for t in th: t.join()
    dt = time.perf_counter() - t0
    print(json.dumps(dict(scene=scene, spp=spp, mode=f"{n_ctx} contexts, stagger {stagger:.2f} frame", ms_per_frame=dt / (n_ctx * frames) * 1e3,
                          mrays=rays * n_ctx * frames / dt * 1e-6, gain=single / (dt / (n_ctx * frames)))), flush=True)
