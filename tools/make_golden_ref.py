"""Golden vectors generated FROM THE REFERENCE ITSELF (oracle/_ref/libref_device.so: the reference's device.cu, disney/*.cuh,
math/random/sample_methods headers and camera.cpp compiled for the host by oracle/ref/Makefile), so that the pin travels to
machines where /root/reference does not exist.

    python tools/make_golden_ref.py        # rewrites tests/golden/ref_*.npz   (needs /root/reference)

Frames: the reference's __raygen__ray_gen per pixel, closest hits from the oracle's finder (OptiX itself is closed source).
glibc libm, -ffp-contract=off: reproducible bit for bit on this image."""
import sys
from pathlib import Path

import numpy as np

R = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(R)); sys.path.insert(0, str(R / "tests")); sys.path.insert(0, str(R / "tools"))
import owl_path_tracer_b200 as pt  # noqa: E402
from checkers import Oracle, Ref  # noqa: E402
from make_golden import bsdf_cases  # noqa: E402
from ref_cases import FRAMES, random_cameras, random_cases  # noqa: E402

G = R / "tests" / "golden"


def main():
    r = Ref()
    # sample_disney: the 960-case grid of round 1 + 4096 random / edge cases
    g_m, g_i, g_wo, g_l, g_r = bsdf_cases()
    m2, i2, wo2, l2, r2 = random_cases(1234, 4096)
    mats = np.concatenate([g_m.view(np.float32).reshape(-1, 17), m2])
    material = np.concatenate([g_i, i2 + g_m.shape[0]]).astype(np.int32)
    wo, lobe, rng = np.concatenate([g_wo, wo2]), np.concatenate([g_l, l2]), np.concatenate([g_r, r2])
    out = r.sample_disney(mats, material, wo, lobe, rng)
    np.savez_compressed(G / "ref_bsdf.npz", mats=mats, material=material, wo=wo, lobe_in=lobe, rng_in=rng,
                        f_bits=out["f"].view(np.uint32), wi_bits=out["wi"].view(np.uint32), pdf_bits=out["pdf"].view(np.uint32),
                        lobe_out=out["lobe"], rng_out=out["rng"])
    # RNG, camera, ONB
    seeds = [(0, 0), (1, 0), (0, 1), (511, 511), (1919, 1079), (7, 1083), (2 ** 31, 5), (2 ** 32 - 1, 2 ** 32 - 1)]
    draws, states = [], []
    for x, y in seeds:
        d, s = r.rng(x, y, 16)
        draws.append(d.view(np.uint32)); states.append(s)
    frm, at, up, fov, wh = random_cameras(11, 64)
    cams = np.stack([r.camera(frm[i], at[i], up[i], float(fov[i]), int(wh[i, 0]), int(wh[i, 1])).view(np.uint32).reshape(-1)
                     for i in range(64)])
    rs = np.random.RandomState(3)
    n = rs.normal(size=(512, 3)).astype(np.float32)
    n /= np.linalg.norm(n, axis=1, keepdims=True)
    n = n.astype(np.float32)
    n[:8] = np.float32(1 / np.sqrt(3))
    w = rs.normal(size=n.shape).astype(np.float32)
    o = r.onb(n, w)
    np.savez_compressed(G / "ref_units.npz", seeds=np.array(seeds, np.uint64), draws_bits=np.array(draws), states=np.array(states, np.uint32),
                        cam_from=frm, cam_at=at, cam_up=up, cam_fov=fov, cam_wh=wh, cam_bits=cams, onb_n=n, onb_w=w,
                        onb_bits=np.stack([o["t"], o["b"], o["local"], o["world"]], 1).view(np.uint32))
    # whole frames
    for name, kw, w_, h_, spp in FRAMES:
        sc = pt.Scene(scene=name, width=w_, height=h_, spp=spp, depth=8, **kw)
        orc = Oracle(det=False).load_scene(sc)
        a = Ref().load_scene(sc).render(orc)
        np.savez_compressed(G / f"ref_frame_{name}.npz", mean_bits=a["mean"].view(np.uint32), rgba=a["rgba"], rays=np.uint64(a["rays"]))
    print("reference golden vectors written to", G)


if __name__ == "__main__":
    main()
