"""b200pt_pipeline_*: the reference's real workload is test_loop — frames back to back with only the materials changing
(application.hpp:89-108).  Frames rendered through the pipeline (frame k+1 starts on a second context while frame k
drains) must be the frames of the blocking call, bit for bit, and equal to the oracle."""
import numpy as np
import pytest

import owl_path_tracer_b200 as pt
from checkers import Oracle

pytestmark = pytest.mark.gpu

SWEEP = (("metallic", 1.0), ("roughness", 0.1), ("clearcoat", 1.0), ("specular_transmission", 1.0), ("sheen", 1.0))


@pytest.mark.parametrize("slots", [1, 2, 3])
def test_pipelined_sweep_frames_equal_blocking_frames_and_the_oracle(slots):
    sc = pt.Scene(scene="mitsuba", width=640, height=360, spp=3, depth=8, env="auto", env_intensity=1.0)   # > 2^17 lanes: parts + parking
    ctxs = [pt.Context() for _ in range(slots)]
    for c in ctxs:
        c.bind(sc)
    mats = []
    for attr, val in SWEEP:
        sc.set_material("outside", attr, val)
        mats.append(sc.materials().copy())
    blocking = [ctxs[0].render_to_host(m).copy() for m in mats]
    pipe = pt.Pipeline(ctxs, start_fraction=0.6)
    for rep in range(2):   # the pipeline object is reusable after wait()
        outs = [np.zeros((sc.height, sc.width), np.uint32) for _ in mats]
        for m, o in zip(mats, outs):
            pipe.submit(m, o)
        pipe.wait()
        for k, (a, b) in enumerate(zip(outs, blocking)):
            assert (a == b).all(), (rep, k)
    pipe.close()
    o = Oracle(det=True).load_scene(sc)
    for k in (0, 3):
        o.set_materials(mats[k])
        assert (blocking[k] == o.render()["rgba"]).all()
    # the contexts are ordinary contexts again
    assert (ctxs[-1].render_to_host(mats[1]) == blocking[1]).all()
    for c in ctxs:
        c.close()


def test_pipeline_reports_a_frame_error():
    sc = pt.Scene(scene="cornell-box", width=64, height=48, spp=2, depth=4)
    a, b = pt.Context(), pt.Context()
    a.bind(sc)
    b.bind(sc)
    pipe = pt.Pipeline([a, b])
    bad = sc.materials()[:1].copy()          # one material for a scene that indexes six
    pipe.submit(sc.materials())
    pipe.submit(bad)
    with pytest.raises(pt.B200ptError):
        pipe.wait()
    pipe.submit(sc.materials())              # still usable
    pipe.wait()
    pipe.close()
    a.close()
    b.close()


def test_path_tracer_pipeline_writes_the_same_files(tmp_path):
    """path_tracer --pipeline 3: the sweep's frames are queued and overlapped; same file names, same bytes"""
    import subprocess
    from checkers import REPO
    exe = REPO / "owl_path_tracer_b200" / "bin" / "path_tracer"
    args = ["--assets", str(REPO / "assets"), "--scene", "mitsuba", "--size", "640", "360", "--spp", "4", "--depth", "6", "--env", "auto",
            "--env-intensity", "1"]
    outs = []
    for extra in ([], ["--pipeline", "3"]):
        d = tmp_path / ("p" if extra else "b")
        d.mkdir()
        r = subprocess.run([str(exe), *args, *extra, "--out", str(d)], capture_output=True, text=True, timeout=300)
        assert r.returncode == 0, r.stderr[-500:]
        outs.append({f.name: f.read_bytes() for f in sorted(d.glob("*.png"))})
    assert len(outs[0]) == 5 and outs[0] == outs[1]
