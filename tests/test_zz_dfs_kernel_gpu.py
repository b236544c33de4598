"""ci_dfs.cu (levels >= 3, lanes over the row's neighbours) against the oracle and against the other CI kernels.

The level plan sends a row to ci_dfs_kernel at l >= 3 when it has 33..63 neighbours, and from level
CUSK_DFS_NARROW_LEVEL (8) on also when CUSK_DFS_MIN_D (20) <= d <= 32 and d - l >= CUSK_DFS_MIN_SPAN (8); the knobs are
read on every call, so these tests lower them to put EVERY row with more than l neighbours of a
small case on the kernel (all levels 3..14, one and two columns per lane, all three modes: individual-level engine,
uniform N, per-pair N with NaN and a time index), and switch the kernel off (CUSK_DFS=0) to compare routes on the same
input: PC-stable results do not depend on which kernel ran a row.
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


class dfs_knobs:
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


ALL_ROWS = dict(CUSK_DFS="1", CUSK_DFS_MIN_SPAN=1, CUSK_DFS_MIN_D=1, CUSK_DFS_NARROW_LEVEL=3)


def test_dfs_all_levels_individual_level_engine(ctx, oracle):
    m, p, Cm = _deep_case()
    Th = oracle.threshold_array(NS, ALPHA)
    ro = oracle.skeleton(Cm, Th, 14, near_band=NEAR_BAND)
    assert ro["l"] == 15 and len(ro["near"]) == 0
    with dfs_knobs(**ALL_ROWS):
        rg = ctx.skeleton(Cm, Th, 14)
    assert rg["l"] == ro["l"]
    assert (rg["G"] == ro["G"]).all()
    assert rg["stats"]["tests_canonical"][:15] == ro["stats"][:15].tolist()
    assert (rg["SepSet"] == ro["SepSet"]).all()
    assert np.allclose(rg["pMax"], ro["pMax"], rtol=1e-5, atol=1e-7)


@pytest.mark.parametrize("general_n", [False, True])
def test_dfs_all_levels_hetcor_engine(ctx, oracle, general_n):
    import cigwas_b200
    from cigwas_b200 import synth

    m, p, Cm = _deep_case()
    n = m + p
    N = (synth.het_sample_sizes(n, m, NS, stream=9, lo=50000, na_rate=0.005) if general_n
         else np.full((n, n), float(NS), np.float32))
    th = cigwas_b200.hetcor_threshold(ALPHA)
    ti = np.zeros(n, np.int32)
    ti[m:] = 1
    ro = oracle.hetcor_skeleton(Cm, np.ones((n, n), np.int32), N, th, 14, ti, near_band=NEAR_BAND)
    with dfs_knobs(**ALL_ROWS):
        rg = ctx.hetcor_skeleton(Cm, None, N, th, 14, ti)
    assert ro["l"] == 15 and rg["l"] == ro["l"]
    _assert_adj_equal(rg["G"], ro["G"], ro["near"])
    assert rg["stats"]["tests_canonical"][:15] == ro["stats"][:15].tolist()


def _trait_heavy_block(m=240, p=6, causal=50, stream=77):
    """traits with `causal` associated markers each: with 50, one trait row keeps 37-38 neighbours (two columns per
    lane) and the others 27-31, at every level"""
    from cigwas_b200 import synth

    return synth.c2_dense_block(m, p, stream=stream, causal_per_trait=causal, beta_lo=0.05, beta_hi=0.1, num_samples=NS)


@pytest.mark.parametrize("maxlevel", [3, 4])
def test_dfs_wide_rows_against_oracle(ctx, oracle, maxlevel):
    m, p = 240, 6
    Cm = _trait_heavy_block(m, p)
    n = m + p
    N = np.full((n, n), float(NS), np.float32)
    ti = np.zeros(n, np.int32)
    ti[m:] = 1
    th = oracle.hetcor_threshold(ALPHA)
    ro = oracle.hetcor_skeleton(Cm, np.ones((n, n), np.int32), N, th, maxlevel, ti, near_band=NEAR_BAND)
    rg = ctx.hetcor_skeleton(Cm, None, N, th, maxlevel, ti)
    assert max(rg["stats"]["max_degree"][3:maxlevel + 1]) >= 32, rg["stats"]["max_degree"]
    _assert_adj_equal(rg["G"], ro["G"], ro["near"])
    assert rg["l"] == ro["l"]
    assert rg["stats"]["tests_canonical"][:maxlevel + 1] == ro["stats"][:maxlevel + 1].tolist()


def test_dfs_route_equals_other_kernels_deep(ctx):
    """levels the oracle cannot reach in test time (10^8+ tests): the same block with ci_dfs on (default knobs) and off
    (rows fall back to ci_small / ci_level) must give the same adjacency and the same canonical counts"""
    import cigwas_b200

    m, p = 240, 6
    Cm = _trait_heavy_block(m, p)
    n = m + p
    N = np.full((n, n), float(NS), np.float32)
    ti = np.zeros(n, np.int32)
    ti[m:] = 1
    th = cigwas_b200.hetcor_threshold(ALPHA)
    ml = 6  # 1.1e8 tests at level 6
    with dfs_knobs(CUSK_DFS="0"):
        off = ctx.hetcor_skeleton(Cm, None, N, th, ml, ti)
    on = ctx.hetcor_skeleton(Cm, None, N, th, ml, ti)
    assert on["l"] == off["l"]
    assert (on["G"] == off["G"]).all()
    assert on["stats"]["tests_canonical"] == off["stats"]["tests_canonical"]
    assert max(on["stats"]["max_degree"][3:ml + 1]) >= 32
    Th = cigwas_b200.threshold_array(NS, ALPHA)
    with dfs_knobs(CUSK_DFS="0"):
        off = ctx.skeleton(Cm, Th, 5)
    on = ctx.skeleton(Cm, Th, 5)
    assert (on["G"] == off["G"]).all() and (on["SepSet"] == off["SepSet"]).all()
    assert np.array_equal(on["pMax"], off["pMax"])
