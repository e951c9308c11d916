"""Freeze known-answer vectors of the oracle into tests/golden/ (the reference ships
no golden vectors, so the oracle's own outputs are the pin).  Deterministic: integer RNG, and the
deterministic-math oracle for everything floating point, so the files reproduce bit for bit.

    python tools/make_golden.py        # rewrites tests/golden/*.npz
"""
import sys
from pathlib import Path

import numpy as np

R = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(R)); sys.path.insert(0, str(R / "tests"))
import owl_path_tracer_b200 as pt  # noqa: E402
from checkers import Oracle  # noqa: E402

G = R / "tests" / "golden"


def bsdf_cases():
    """A grid of (material, wo, previous lobe, rng state) covering every lobe and branch."""
    base = dict(base_color_r=.8, base_color_g=.6, base_color_b=.3, subsurface=0, metallic=0, specular=.5, specular_tint=.3,
                roughness=.5, anisotropic=0, sheen=0, sheen_tint=.5, clearcoat=0, clearcoat_gloss=.5, ior=1.5,
                specular_transmission=0, specular_transmission_roughness=0, emission=0)
    variants = [
        {}, dict(roughness=1.0), dict(roughness=0.05), dict(sheen=1.0), dict(sheen=.7, sheen_tint=1.0),
        dict(metallic=1.0), dict(metallic=1.0, roughness=.2), dict(metallic=1.0, roughness=.1, anisotropic=1.0),
        dict(metallic=1.0, roughness=.6, anisotropic=.5, specular_tint=1.0), dict(metallic=.5, roughness=.2),
        dict(clearcoat=1.0, clearcoat_gloss=.9), dict(clearcoat=1.0, clearcoat_gloss=.1), dict(clearcoat=.5, metallic=.3),
        dict(specular_transmission=1.0, specular_transmission_roughness=.01), dict(specular_transmission=1.0, ior=1.33),
        dict(specular_transmission=1.0, specular_transmission_roughness=.4, roughness=.3),
        dict(specular_transmission=.5, metallic=.2, clearcoat=.4, sheen=.3),
        dict(specular_transmission=1.0, ior=1.0), dict(base_color_r=0, base_color_g=0, base_color_b=0, sheen=1.0),
        dict(metallic=0, specular_transmission=0, clearcoat=0, roughness=0.0),
    ]
    mats = np.zeros(len(variants), pt.MATERIAL_DTYPE)
    for i, v in enumerate(variants):
        for k, x in {**base, **v}.items():
            mats[i][k] = x
    rs = np.random.RandomState(0x5EED)
    n_per = 48
    material = np.repeat(np.arange(len(variants), dtype=np.int32), n_per)
    n = material.size
    wo = rs.normal(size=(n, 3)).astype(np.float32)
    wo /= np.linalg.norm(wo, axis=1, keepdims=True)
    wo[::7, 2] = np.abs(wo[::7, 2])           # more front-facing cases
    wo[3::16] = np.array([0, 0, 1], np.float32)  # normal incidence
    wo[5::16, 2] *= 1e-3                         # grazing
    wo /= np.linalg.norm(wo, axis=1, keepdims=True)
    lobe = rs.randint(-1, 4, size=n).astype(np.int32)
    rng = rs.randint(0, 2 ** 32, size=n, dtype=np.uint64).astype(np.uint32)
    return mats, material, wo.astype(np.float32), lobe, rng


def main():
    G.mkdir(parents=True, exist_ok=True)
    o = Oracle(det=True)
    # RNG: first 8 draws of a few pixels (random.hpp:34-85)
    seeds = [(0, 0), (1, 0), (0, 1), (511, 511), (1919, 1079), (7, 1080 + 3)]
    draws, states = [], []
    for x, y in seeds:
        d, s = o.rng(x, y, 8)
        draws.append(d.view(np.uint32)); states.append(s)
    np.savez(G / "rng.npz", seeds=np.array(seeds, np.uint32), draws_bits=np.array(draws), states=np.array(states, np.uint32))
    # BSDF
    mats, material, wo, lobe, rng = bsdf_cases()
    r = o.sample_disney(mats, material, wo, lobe, rng)
    np.savez(G / "bsdf.npz", mats=mats.view(np.float32).reshape(-1, 17), material=material, wo=wo, lobe_in=lobe, rng_in=rng,
             f_bits=r["f"].view(np.uint32), wi_bits=r["wi"].view(np.uint32), pdf_bits=r["pdf"].view(np.uint32),
             lobe_out=r["lobe"], rng_out=r["rng"])
    # cornell box: primary-hit map and a 32x32 / 16 spp float image (camera per assets/cornell-box.json)
    sc = pt.Scene(scene="cornell-box", width=64, height=64, spp=16, depth=8, env="color", env_intensity=0.0)
    o2 = Oracle(det=True).load_scene(sc)
    h = o2.primary_hits()
    img = o2.render()
    np.savez_compressed(G / "cornell_64.npz", mesh=h["mesh"], prim=h["prim"], t_bits=h["t"].view(np.uint32),
                        sum_bits=img["sum"].view(np.uint32), rgba=img["rgba"], rays=np.uint64(img["rays"]))
    # cube with auto sky + texture stand-in, sphere with environment stand-in
    for name, kw in (("cube", dict(env="auto", env_intensity=1.0)), ("sphere", dict(env="map", env_intensity=1.0))):
        sc = pt.Scene(scene=name, width=48, height=48, spp=8, depth=8, **kw)
        oo = Oracle(det=True).load_scene(sc)
        img = oo.render()
        np.savez_compressed(G / f"{name}_48.npz", sum_bits=img["sum"].view(np.uint32), rgba=img["rgba"], rays=np.uint64(img["rays"]))
    cam = o.camera([3, 1, 0], [0, 1, 0], [0, 1, 0], 50.0, 512, 512)
    np.savez(G / "camera.npz", cam_bits=cam.view(np.uint32))
    print("golden vectors written to", G)


if __name__ == "__main__":
    main()
