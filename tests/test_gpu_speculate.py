"""GPU: builds that take the previous build's counts instead of waiting for the device ("gpu":{"speculate":1}, the default)
give the bytes a waiting build gives, say so in their counters, and fall back to a waiting build when the counts differ."""
import json

import numpy as np
import pytest

from implisolid_b200 import shapes

pytestmark = pytest.mark.gpu

OB02 = dict(iters=1, c=1.0, projection=1, qem=1)


def _build(lib, sj, mc):
    lib.build_geometry(sj, mc)
    out = (lib.get_v().copy(), lib.get_f().copy(), lib.last_counters(), lib.average_edge())
    lib.finish_geometry()
    return out


def _same(a, b):
    keys = ("cut_cubes", "projection_by_type", "projection_failed", "bisection_capped", "bisected")
    return (a[0].tobytes() == b[0].tobytes() and a[1].tobytes() == b[1].tobytes() and all(a[2][k] == b[2][k] for k in keys) and
            np.float32(a[3]).tobytes() == np.float32(b[3]).tobytes())


@pytest.mark.parametrize("name,shape,R,half,kw", [
    ("chain_ob02", shapes.c4_deep_chain, 96, 1.0, OB02),
    ("chain_two_repeats", shapes.c4_deep_chain, 64, 1.0, dict(OB02, repeats=2)),
    ("cube_cyl_ob02", shapes.c2_cube_minus_cylinder, 80, 0.8, OB02),
    ("metaballs_mc_only", shapes.c3_metaballs, 64, 1.2, {}),
    ("sphere_serial_sum", shapes.c1_sphere, 40, 0.55, dict(OB02, exact_reduce=1)),
    ("cube_cyl_subdivision", shapes.c2_cube_minus_cylinder, 40, 0.8, dict(OB02, subdiv=1)),
])
def test_speculative_builds_change_nothing(gpu_lib, name, shape, R, half, kw):
    sj = shape()
    kw = dict(kw)
    subdiv = kw.pop("subdiv", 0)

    def settings(**more):
        mc = json.loads(shapes.mc_settings(R, half, **kw, **more))
        if subdiv:      # with noise: the glibc rand() stream runs on from build to build, so every build compared starts it afresh
            mc["subdiv"] = {"enabled": 1}
            mc["debug"] = {"post_subdiv_noise": 0.001}
        return json.dumps(mc)

    waiting, spec = settings(speculate=0), settings()
    gpu_lib.spoil_speculation(0)
    gpu_lib.srand(1)
    ref = _build(gpu_lib, sj, waiting)
    assert ref[2]["speculative"] == 0
    runs = []
    for i in range(3):
        gpu_lib.srand(1)
        runs.append(_build(gpu_lib, sj, spec))
    assert all(r[2]["speculative"] == 1 for r in runs)      # the waiting build left its counts behind
    assert all(r[2]["speculation_redone"] == 0 for r in runs)
    for r in runs:
        assert _same(ref, r), name


def test_another_spec_is_not_taken_for_the_previous_one(gpu_lib):
    a, b = shapes.c4_deep_chain(), shapes.c2_cube_minus_cylinder()
    mc_a, mc_b = shapes.mc_settings(64, 1.0, **OB02), shapes.mc_settings(64, 0.8, **OB02)
    gpu_lib.spoil_speculation(0)
    first = _build(gpu_lib, a, mc_a)
    other = _build(gpu_lib, b, mc_b)
    assert other[2]["speculative"] == 0
    again = _build(gpu_lib, a, mc_a)
    assert again[2]["speculative"] == 0             # one spec is remembered, the last one
    assert _same(first, again)
    finer = _build(gpu_lib, a, shapes.mc_settings(72, 1.0, **OB02))
    assert finer[2]["speculative"] == 0


def test_wrong_counts_are_caught_and_the_build_is_done_again(gpu_lib, capfd):
    sj, mc = shapes.c4_deep_chain(), shapes.mc_settings(96, 1.0, **OB02)
    gpu_lib.spoil_speculation(0)
    ref = _build(gpu_lib, sj, mc)
    ok = _build(gpu_lib, sj, mc)
    assert ok[2]["speculative"] == 1 and _same(ref, ok)
    gpu_lib.spoil_speculation(1)
    redone = _build(gpu_lib, sj, mc)
    assert redone[2]["speculation_redone"] == 1 and redone[2]["speculative"] == 0
    assert _same(ref, redone)
    assert "building again" in capfd.readouterr().err
    after = _build(gpu_lib, sj, mc)                 # the build that was done again left good counts behind
    assert after[2]["speculative"] == 1 and after[2]["speculation_redone"] == 0 and _same(ref, after)
