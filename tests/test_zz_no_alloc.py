"""b200pt_trace_static (include/b200pt.h): the no-allocation variant of b200pt_trace
(no counterpart in the reference).  Collected last on purpose.  Through the C ABI, two calls in a row on the
shared static scratch, each against the existing entry point: bit-equal records (the tolerance of the other hit tests)."""
import numpy as np
import pytest

import owl_path_tracer_b200 as pt

pytestmark = pytest.mark.gpu


def rays(seed, n):
    rs = np.random.RandomState(seed)
    o = (np.array([0.2, 1.0, 0.0], np.float32) + rs.uniform(-.6, .6, (n, 3))).astype(np.float32)
    d = rs.normal(size=(n, 3)).astype(np.float32)
    d /= np.linalg.norm(d, axis=1, keepdims=True)
    return o, d.astype(np.float32)


def same(a, b):
    return all((np.ascontiguousarray(a[k]).view(np.uint32) == np.ascontiguousarray(b[k]).view(np.uint32)).all()
               for k in ("mesh", "prim", "t", "u", "v"))


def test_trace_static_equals_trace_on_two_calls_in_a_row():
    sc = pt.Scene(scene="cornell-box", width=8, height=8, spp=1, depth=1)
    ctx = pt.Context()
    ctx.bind(sc)
    # second batch is smaller than the first: stale records of call 1 stay in the scratch behind it and must not leak
    for seed, n in ((11, 50000), (12, 7777)):
        o, d = rays(seed, n)
        want = ctx.trace(o, d)
        got = ctx.trace_static(o, d)
        assert same(got, want)
        assert (want["mesh"] >= 0).sum() > n // 4
    o, d = rays(13, 65536)
    assert same(ctx.trace_static(o, d), ctx.trace(o, d))            # exactly the maximum
    o, d = rays(14, 65537)
    with pytest.raises(pt.B200ptError, match="exceed the fixed scratch of 65536"):
        ctx.trace_static(o, d)
    ctx.close()
