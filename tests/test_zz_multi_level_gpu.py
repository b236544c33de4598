"""Self-planned deep levels (ci_small.cu, ci_small_multi_kernel): levels 4..14 of a light het block inside three cooperative
launches (degrees, adjacency and work items of a level built in the launch, grid barriers between the phases) against the
oracle and against the launch-per-level route (CUSK_SELF_PLAN=0) on the same input.  PC-stable: the adjacency, the canonical
test counts and the per-level maximum degrees do not depend on the route.
"""
import os

import numpy as np
import pytest

from tests.test_skeleton_gpu import ALPHA, NEAR_BAND, NS, _assert_adj_equal, _deep_case

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def ctx():
    import cigwas_b200

    c = cigwas_b200.Context(0)
    yield c
    c.close()


class env:
    def __init__(self, **kw):
        self.kw = {k: str(v) for k, v in kw.items()}

    def __enter__(self):
        self.saved = {k: os.environ.get(k) for k in self.kw}
        os.environ.update(self.kw)

    def __exit__(self, *a):
        for k, v in self.saved.items():
            if v is None:
                os.environ.pop(k, None)
            else:
                os.environ[k] = v


def _inputs(general_n):
    import cigwas_b200
    from cigwas_b200 import synth

    m, p, Cm = _deep_case()
    n = m + p
    N = (synth.het_sample_sizes(n, m, NS, stream=9, lo=50000, na_rate=0.005) if general_n
         else np.full((n, n), float(NS), np.float32))
    ti = np.zeros(n, np.int32)
    ti[m:] = 1
    return Cm, N, cigwas_b200.hetcor_threshold(ALPHA), ti


@pytest.mark.parametrize("general_n", [False, True])
@pytest.mark.parametrize("maxlevel", [14, 9, 5])
def test_multi_level_route_equals_level_per_launch_route(ctx, oracle, general_n, maxlevel):
    Cm, N, th, ti = _inputs(general_n)
    n = Cm.shape[0]
    with env(CUSK_SELF_PLAN=0):
        r0 = ctx.hetcor_skeleton(Cm, None, N, th, maxlevel, ti)
    with env(CUSK_SELF_PLAN=1):
        r1 = ctx.hetcor_skeleton(Cm, None, N, th, maxlevel, ti)
        r2 = ctx.hetcor_skeleton(Cm, None, N, th, maxlevel, ti)  # a second call: the control blocks start from zero again
    with env(CUSK_SELF_PLAN=1, CUSK_SELF_COOP=0):  # the same kernels as plain (programmatic dependent) launches
        r3 = ctx.hetcor_skeleton(Cm, None, N, th, maxlevel, ti)
    assert r3["l"] == r0["l"] and (r3["G"] == r0["G"]).all()
    assert r3["stats"]["tests_canonical"] == r0["stats"]["tests_canonical"]
    s0, s1 = r0["stats"], r1["stats"]
    assert s0["max_degree"][3] <= 22, "the case must qualify for the self-planned route"
    assert r1["l"] == r0["l"] == r2["l"]
    assert (r1["G"] == r0["G"]).all() and (r2["G"] == r0["G"]).all()
    assert s1["tests_canonical"] == s0["tests_canonical"]
    assert s1["max_degree"] == s0["max_degree"]
    assert s1["edges_after"] == s0["edges_after"]
    # levels 4..maxlevel: at most three launches instead of (at least) two per level
    deep = maxlevel - 3
    assert s1["kernel_launches"] <= s0["kernel_launches"] - 2 * deep + 3, (s0["kernel_launches"], s1["kernel_launches"])
    assert s1["kernel_launches"] < s0["kernel_launches"]
    ro = oracle.hetcor_skeleton(Cm, np.ones((n, n), np.int32), N, th, maxlevel, ti, near_band=NEAR_BAND)
    assert r1["l"] == ro["l"]
    _assert_adj_equal(r1["G"], ro["G"], ro["near"])
    assert s1["tests_canonical"][:maxlevel + 1] == ro["stats"][:maxlevel + 1].tolist()


def test_multi_level_route_reports_level_times(ctx):
    Cm, N, th, ti = _inputs(False)
    ctx.set_level_timing(True)
    try:
        with env(CUSK_SELF_PLAN=1):
            r = ctx.hetcor_skeleton(Cm, None, N, th, 14, ti)
    finally:
        ctx.set_level_timing(False)
    st = r["stats"]
    for l in range(1, 15):
        assert st["level_ms"][l] > 0.0, (l, st["level_ms"])
        assert 0.0 < st["ci_kernel_ms"][l] <= st["level_ms"][l] + 1e-3, (l, st["ci_kernel_ms"], st["level_ms"])


def test_multi_level_route_ends_at_the_first_empty_level(ctx, oracle):
    """a sparse graph: no row keeps more than a handful of neighbours, so the launch that holds level 4 finds it (or a
    later level) empty and marks the rest of the call done, as the level-per-launch route does"""
    import cigwas_b200
    from cigwas_b200 import synth

    Cm = synth.sem_correlation(300, 5, stream=21, num_samples=NS, causal_per_trait=6)
    n = Cm.shape[0]
    N = np.full((n, n), float(NS), np.float32)
    th = cigwas_b200.hetcor_threshold(ALPHA)
    with env(CUSK_SELF_PLAN=0):
        r0 = ctx.hetcor_skeleton(Cm, None, N, th, 14, None)
    with env(CUSK_SELF_PLAN=1):
        r1 = ctx.hetcor_skeleton(Cm, None, N, th, 14, None)
    assert r1["l"] == r0["l"]
    assert (r1["G"] == r0["G"]).all()
    assert r1["stats"]["tests_canonical"] == r0["stats"]["tests_canonical"]
    assert r1["stats"]["levels_run"] == r0["stats"]["levels_run"]
    ro = oracle.hetcor_skeleton(Cm, np.ones((n, n), np.int32), N, th, 14, np.zeros(n, np.int32), near_band=NEAR_BAND)
    assert r1["l"] == ro["l"]
    _assert_adj_equal(r1["G"], ro["G"], ro["near"])
