import os, sys, json
sys.path.insert(0, '/root/repo'); sys.path.insert(0, '/root/repo/tests')
os.environ["B200PT_QUIET"]="1"
import owl_path_tracer_b200 as pt
for (w,h,spp) in ((1920,1080,128),(3840,2160,32),(7680,4320,8),(960,540,512)):
    sc = pt.Scene(scene="dragon", width=w, height=h, spp=spp, depth=8, env="color", env_intensity=0.0)
    ctx = pt.Context(flags=0); ctx.bind(sc); ctx.render(); ctx.render()
    st = ctx.stats()
    print(json.dumps(dict(size=[w,h], spp=spp, render_ms=st['render_ms'], mrays=st['rays']/st['render_ms']*1e-3, rays=st['rays'], launches=st['kernel_launches'])), flush=True)
    ctx.close()
