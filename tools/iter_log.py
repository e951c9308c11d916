"""Per-iteration log (queue length, parked lanes, extend ms, shade ms) of one dragon frame.  Needs a development build
(-DB200PT_DEV=1, tools/build_variants.sh):   python tools/iter_log.py <libb200pt_dev.so> <out.txt> [spp]"""
import os, sys
from pathlib import Path
R = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(R)); sys.path.insert(0, str(R / "tests"))
os.environ["B200PT_QUIET"] = "1"
import owl_path_tracer_b200 as pt
pt.use_library(sys.argv[1])
sc = pt.Scene(scene="dragon", width=1920, height=1080, spp=int(sys.argv[3]) if len(sys.argv) > 3 else 32, depth=8, env="color", env_intensity=0.0)
ctx = pt.Context(flags=pt.FLAG_TIMING); ctx.bind(sc); ctx.render()
os.environ["B200PT_ITER_LOG"] = sys.argv[2]
ctx.render()
print(ctx.stats()["render_ms"], ctx.stats()["iterations"])
