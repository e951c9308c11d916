"""Quick GPU-vs-oracle check + timing (development helper; the real gates are tests/ -m gpu)."""
import sys, time, json
from pathlib import Path
R = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(R)); sys.path.insert(0, str(R / "tests"))
import numpy as np
import owl_path_tracer_b200 as pt
from checkers import Oracle, rel_mse

def bits(a): return np.ascontiguousarray(a).view(np.uint32)

def check(scene, w, h, spp, depth, **kw):
    sc = pt.Scene(scene=scene, width=w, height=h, spp=spp, depth=depth, **kw)
    ctx = pt.Context(flags=pt.FLAG_COUNTERS)
    ctx.bind(sc)
    o = Oracle(det=True).load_scene(sc)
    hg, ho = ctx.primary_hits(), o.primary_hits()
    same = (hg['mesh']==ho['mesh'])&(hg['prim']==ho['prim'])&(bits(hg['t'])==bits(ho['t']))&(bits(hg['u'])==bits(ho['u']))&(bits(hg['v'])==bits(ho['v']))
    ctx.render()
    st = ctx.stats()
    t=time.time(); ro = o.render(); to=time.time()-t
    g = ctx.framebuffer_f32(); rg = ctx.framebuffer_rgba8()
    print(json.dumps(dict(scene=scene, size=[w,h], spp=spp, tris=sc.n_triangles, primary_mismatch=int((~same).sum()),
        mean_bit_mismatch=int((bits(g[...,:3])!=bits(ro['mean'][...,:3])).sum()), rgba_mismatch=int((rg!=ro['rgba']).sum()),
        relmse=rel_mse(g, ro['mean']), rays_gpu=st['rays'], rays_oracle=ro['rays'], render_ms=st['render_ms'],
        mrays=st['rays']/st['render_ms']*1e-3, oracle_s=to, nodes_per_ray=st['nodes_visited']/max(1,st['rays']),
        tris_per_ray=st['tris_tested']/max(1,st['rays']), build_ms=st['build_ms'], bvh_nodes=st['bvh_nodes'], iters=st['iterations'])))
    ctx.close()

def timing(scene, w, h, spp, depth, flags=0, **kw):
    sc = pt.Scene(scene=scene, width=w, height=h, spp=spp, depth=depth, **kw)
    ctx = pt.Context(flags=flags)
    ctx.bind(sc)
    ctx.render(); ctx.render()
    st = ctx.stats()
    print(json.dumps(dict(scene=scene, size=[w,h], spp=spp, tris=sc.n_triangles, render_ms=st['render_ms'], extend_ms=st['extend_ms'], shade_ms=st['shade_ms'],
        mrays=st['rays']/st['render_ms']*1e-3, rays=st['rays'], iters=st['iterations'], build_ms=st['build_ms'], bvh_nodes=st['bvh_nodes'], launches=st['kernel_launches'])))
    ctx.close()

if __name__ == "__main__":
    import os
    os.environ.setdefault("B200PT_QUIET", "1")
    if len(sys.argv) > 1 and sys.argv[1] == "ab":
        timing("dragon", 1920, 1080, 32, 8, flags=pt.FLAG_TIMING, env="color", env_intensity=0.0)
        timing("car", 1920, 1080, 32, 8, flags=pt.FLAG_TIMING, env="color", env_intensity=0.0)
        sys.exit(0)
    if len(sys.argv) > 1 and sys.argv[1] == "ab2":
        # product configuration (two concurrent sub-frames, no per-kernel events), parity at a size where walks get parked
        check("dragon", 640, 360, 4, 8, env="color", env_intensity=0.0)
        timing("dragon", 1920, 1080, 128, 8, env="color", env_intensity=0.0)
        timing("car", 1920, 1080, 128, 8, env="color", env_intensity=0.0)
        sys.exit(0)
    if len(sys.argv) > 1 and sys.argv[1] == "parts":
        # concurrent sub-frames: frame time without per-kernel events, for B200PT_PARTS = 1..4
        for scene, spp in (("dragon", 32), ("car", 32), ("dragon", 128)):
            for parts in ("1", "2", "3", "4"):
                os.environ["B200PT_PARTS"] = parts
                sc = pt.Scene(scene=scene, width=1920, height=1080, spp=spp, depth=8, env="color", env_intensity=0.0)
                ctx = pt.Context(flags=0); ctx.bind(sc); ctx.render(); ctx.render()
                st = ctx.stats()
                print(json.dumps(dict(scene=scene, spp=spp, parts=int(parts), render_ms=st['render_ms'],
                                      mrays=st['rays'] / st['render_ms'] * 1e-3, rays=st['rays'], launches=st['kernel_launches'])), flush=True)
                ctx.close()
        os.environ.pop("B200PT_PARTS", None)
        sys.exit(0)
    if len(sys.argv) > 1 and sys.argv[1] == "streams":
        # several RNG streams per pixel: frame time at equal total work
        for scene, spp in (("dragon", 128), ("car", 128)):
            for V in (1, 2, 4, 8, 16):
                sc = pt.Scene(scene=scene, width=1920, height=1080, spp=spp, depth=8, env="color", env_intensity=0.0)
                ctx = pt.Context(flags=0, streams=V); ctx.bind(sc); ctx.render(); ctx.render()
                st = ctx.stats()
                print(json.dumps(dict(scene=scene, spp=spp, streams=V, render_ms=st['render_ms'], mrays=st['rays'] / st['render_ms'] * 1e-3,
                                      rays=st['rays'], paths=st['paths'], launches=st['kernel_launches'])), flush=True)
                ctx.close()
        sys.exit(0)
    if len(sys.argv) > 1 and sys.argv[1] == "nee":
        # NEE mode: throughput with the shadow kernel, and error at equal spp / equal time against a long render
        for scene, w, h, spp, env in (("cornell-box", 512, 512, 64, "color"), ("sphere", 1024, 1024, 32, "map"), ("sphere", 1024, 1024, 32, "float"),
                                      ("car", 1920, 1080, 32, "color"), ("dragon", 1920, 1080, 32, "color")):
            is_float = env == "float"
            env = "map" if is_float else env
            sc = pt.Scene(scene=scene, width=w, height=h, spp=spp, depth=8, env=env, env_intensity=1.0 if env == "map" else 0.0, env_float=is_float)
            res = {}
            for tag, fl in (("ref", pt.FLAG_TIMING), ("nee", pt.FLAG_TIMING | pt.FLAG_NEE)):
                ctx = pt.Context(flags=fl); ctx.bind(sc); ctx.render(); ctx.render()
                st = ctx.stats(); res[tag] = (st, ctx.framebuffer_f32()[..., :3].copy()); ctx.close()
            scl = pt.Scene(scene=scene, width=w, height=h, spp=spp * 16, depth=8, env=env, env_intensity=1.0 if env == "map" else 0.0, env_float=is_float)
            ctx = pt.Context(flags=pt.FLAG_NEE); ctx.bind(scl); ctx.render(); truth = ctx.framebuffer_f32()[..., :3].copy(); ctx.close()
            out = dict(scene=scene + ("+float env" if is_float else ""), size=[w, h], spp=spp)
            for tag in ("ref", "nee"):
                st, img = res[tag]
                out[tag] = dict(render_ms=st['render_ms'], extend_ms=st['extend_ms'], shade_ms=st['shade_ms'], shadow_ms=st['shadow_ms'],
                                rays=st['rays'], shadow_rays=st['shadow_rays'], n_lights=st['n_lights'],
                                mrays_incl_shadow=(st['rays'] + st['shadow_rays']) / st['render_ms'] * 1e-3,
                                relmse_vs_16x_nee=rel_mse(img, truth), mean=float(img.mean()))
            out["truth_mean"] = float(truth.mean())
            print(json.dumps(out), flush=True)
        sys.exit(0)
    if len(sys.argv) > 1 and sys.argv[1] == "timing":
        timing("cornell-box", 512, 512, 64, 8, flags=pt.FLAG_TIMING, env="color", env_intensity=0.0)
        timing("dragon", 1920, 1080, 16, 8, flags=pt.FLAG_TIMING, env="color", env_intensity=0.0)
        timing("car", 1920, 1080, 16, 8, flags=pt.FLAG_TIMING, env="color", env_intensity=0.0)
        sys.exit(0)
    check("cornell-box", 128, 128, 8, 8, env="color", env_intensity=0.0)
    check("cube", 128, 128, 8, 8, env="auto", env_intensity=1.0)
    check("sphere", 128, 128, 8, 8, env="map", env_intensity=1.0)
    check("mitsuba", 160, 90, 4, 8, env="auto", env_intensity=1.0)
    check("car", 160, 90, 4, 8, env="color", env_intensity=0.0)
    check("dragon", 160, 90, 4, 8, env="color", env_intensity=0.0)
    timing("cornell-box", 512, 512, 64, 8, env="color", env_intensity=0.0)
    timing("cornell-box", 512, 512, 64, 8, flags=pt.FLAG_TIMING, env="color", env_intensity=0.0)
    timing("dragon", 1920, 1080, 16, 8, env="color", env_intensity=0.0)
    timing("dragon", 1920, 1080, 16, 8, flags=pt.FLAG_TIMING, env="color", env_intensity=0.0)
