"""The deterministic elementary functions (owl_path_tracer_b200/csrc/pt_detmath.h) against numpy
float64, and against the glibc-built oracle.  Tolerance: 1 ulp of binary32 (the functions round a binary64 evaluation
once).  The binary32 build option (-DPT_TRIG_F32=1, measured 1.5 - 3.4 ulp) is not what the product ships."""
import numpy as np
import pytest

from checkers import Emu, Oracle


def ulp_err(got, ref64):
    ref32 = ref64.astype(np.float32)
    ok = np.isfinite(ref32) & (ref32 != 0)
    spacing = np.spacing(np.abs(ref32[ok])).astype(np.float64)
    return np.abs(got[ok].astype(np.float64) - ref64[ok]) / spacing


def grid(lo, hi, n, seed):
    rs = np.random.RandomState(seed)
    return np.concatenate([np.linspace(lo, hi, n), rs.uniform(lo, hi, n)]).astype(np.float32)


CASES = [
    ("sin", lambda x: np.sin(x), np.concatenate([grid(-20, 20, 40000, 1), grid(-100, 100, 40000, 21), grid(-3e4, 3e4, 10000, 22)])),
    ("cos", lambda x: np.cos(x), np.concatenate([grid(-20, 20, 40000, 2), grid(-100, 100, 40000, 23), grid(-3e4, 3e4, 10000, 24)])),
    ("tan", lambda x: np.tan(x), grid(0.0, 6.7, 40000, 3)),      # range of 2*pi*u + 1/pi
    ("atan", lambda x: np.arctan(x), np.concatenate([grid(-50, 50, 30000, 4), grid(-1e6, 1e6, 5000, 5)])),
    ("asin", lambda x: np.arcsin(x), grid(-1, 1, 40000, 6)),
    ("log", lambda x: np.log(x), np.concatenate([grid(1e-7, 1e-1, 30000, 7), grid(0.1, 100, 10000, 8)])),
]
#          max ulp, mean ulp
BOUND = {k: (1.0, 0.26) for k in ("sin", "cos", "tan", "atan", "asin", "log")}


@pytest.mark.parametrize("name,ref,x", CASES, ids=[c[0] for c in CASES])
def test_unary_accuracy(name, ref, x):
    got = Emu().math(name, x)
    e = ulp_err(got, ref(x.astype(np.float64)))
    assert e.max() <= BOUND[name][0], f"{name}: {e.max()} ulp"
    assert e.mean() <= BOUND[name][1], f"{name}: mean {e.mean()} ulp"
    assert np.mean(e <= 0.5000001) > 0.999  # almost always correctly rounded


def test_atan2_and_pow():
    rs = np.random.RandomState(9)
    y = rs.normal(size=60000).astype(np.float32)
    x = rs.normal(size=60000).astype(np.float32)
    e = ulp_err(Emu().math("atan2", y, x), np.arctan2(y.astype(np.float64), x.astype(np.float64)))
    assert e.max() <= 1.0
    b = np.concatenate([rs.uniform(0, 1, 30000), rs.uniform(1e-6, 1e-2, 30000)]).astype(np.float32)
    p = np.concatenate([np.full(30000, 2.2), rs.uniform(0, 1, 30000)]).astype(np.float32)
    e = ulp_err(Emu().math("pow", b, p), np.power(b.astype(np.float64), p.astype(np.float64)))
    assert e.max() <= 1.0


def test_special_values():
    m = Emu().math
    inf, nan = np.float32(np.inf), np.float32(np.nan)
    assert np.isnan(m("sin", [inf, nan])).all() and np.isnan(m("cos", [-inf])).all()
    assert m("atan", [inf])[0] == np.float32(np.pi / 2) and m("atan", [-inf])[0] == np.float32(-np.pi / 2)
    assert np.isnan(m("asin", [1.0000001, -2.0])).all()
    assert m("asin", [1.0])[0] == np.float32(np.pi / 2) and m("asin", [-1.0])[0] == np.float32(-np.pi / 2)
    assert m("log", [0.0])[0] == -inf and np.isnan(m("log", [-1.0]))[0] and m("log", [1.0])[0] == 0.0
    assert m("pow", [0.0, 5.0, 1.0, 0.0], [2.2, 0.0, 7.0, 0.0]).tolist() == [0.0, 1.0, 1.0, 1.0]
    a = m("atan2", [0.0, 0.0, 1.0, -1.0, 0.0], [1.0, -1.0, 0.0, 0.0, 0.0])
    assert a.tolist() == [0.0, np.float32(np.pi), np.float32(np.pi / 2), np.float32(-np.pi / 2), 0.0]


def test_det_oracle_uses_the_same_functions_and_libm_is_close():
    x = grid(-7, 7, 5000, 11)
    e, od, ol = Emu(), Oracle(det=True), Oracle(det=False)
    for fn in ("sin", "cos", "atan"):
        a, b, c = e.math(fn, x), od.math(fn, x), ol.math(fn, x)
        assert (a.view(np.uint32) == b.view(np.uint32)).all()
        assert np.max(np.abs(a.astype(np.float64) - c) / np.spacing(np.abs(c)).astype(np.float64)) <= 2.0
