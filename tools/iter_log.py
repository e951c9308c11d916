import os, sys
sys.path.insert(0, '/root/repo'); sys.path.insert(0, '/root/repo/tests')
os.environ["B200PT_QUIET"]="1"
import owl_path_tracer_b200 as pt
sc = pt.Scene(scene="dragon", width=1920, height=1080, spp=int(os.environ.get("SPP", "32")), depth=8, env="color", env_intensity=0.0)
ctx = pt.Context(flags=pt.FLAG_TIMING); ctx.bind(sc); ctx.render()
os.environ["B200PT_ITER_LOG"] = sys.argv[1]
ctx.render()
print(ctx.stats()["render_ms"], ctx.stats()["iterations"])
