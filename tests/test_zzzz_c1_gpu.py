"""BASELINE.json configs[0] on the GPU: the simulate_dag small DAG (1 600 markers + 10 traits, n = 10 000, alpha 1e-4,
max level 3; cigwas_b200.synth.simulate_dag restates simulation/simulate_dag.R) through both engines of the C ABI
against the oracle.  Same bar as tests/test_skeleton_gpu.py: adjacency identical outside the 1e-9 band, canonical
test counts exact, separating sets identical when no edge sits in the band."""
import numpy as np
import pytest

from tests.test_skeleton_gpu import NEAR_BAND, _assert_adj_equal

pytestmark = pytest.mark.gpu
ALPHA, NS, SNP, TR = 1e-4, 10000, 1600, 10


@pytest.fixture(scope="module")
def ctx():
    import cigwas_b200

    c = cigwas_b200.Context(0)
    yield c
    c.close()


@pytest.fixture(scope="module")
def c1():
    from cigwas_b200 import synth

    return synth.simulate_dag(n=NS, snp=SNP, tr=TR)[0]


@pytest.mark.parametrize("maxlevel", [0, 1, 3, 14])
def test_c1_summary_statistic_engine(ctx, oracle, c1, maxlevel):
    n = c1.shape[0]
    th = oracle.hetcor_threshold(ALPHA)
    N = np.full((n, n), float(NS), np.float32)
    ti = np.zeros(n, np.int32)
    ti[SNP:] = 1
    ones = np.ones((n, n), np.int32)
    ro = oracle.hetcor_skeleton(c1, ones, N, th, maxlevel, ti, near_band=NEAR_BAND)
    rg = ctx.hetcor_skeleton(c1, ones, N, th, maxlevel, ti)
    _assert_adj_equal(rg["G"], ro["G"], ro["near"])
    assert rg["l"] == ro["l"]
    assert rg["stats"]["tests_canonical"][: ro["l"] + 1] == ro["stats"][: ro["l"] + 1].tolist()


def test_c1_individual_level_engine(ctx, oracle, c1):
    Th = oracle.threshold_array(NS, ALPHA)
    ro = oracle.skeleton(c1, Th, 3, near_band=NEAR_BAND)
    rg = ctx.skeleton(c1, Th, 3)
    ndiff = _assert_adj_equal(rg["G"], ro["G"], ro["near"])
    assert rg["l"] == ro["l"]
    assert rg["stats"]["tests_canonical"][: ro["l"] + 1] == ro["stats"][: ro["l"] + 1].tolist()
    if ndiff == 0 and len(ro["near"]) == 0:
        assert (rg["SepSet"] == ro["SepSet"]).all()
        assert np.allclose(rg["pMax"], ro["pMax"], rtol=1e-5, atol=1e-7)
