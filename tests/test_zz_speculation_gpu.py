"""Host-buffer het call (cusk_hetcor_skeleton): the engine starts on the assumption that G is all ones and N one constant
while a helper thread scans the two matrices beside it (api.cu).  Cases: the assumption holds (G not passed; G passed as an
array of ones), and a glance at the first entries passes but the full scan does not -- one late zero in G, one late odd
value in N -- so the call runs a second time with the real inputs.  Every result equals the oracle's."""
import numpy as np
import pytest

from tests.test_skeleton_gpu import ALPHA, NEAR_BAND, NS, _assert_adj_equal

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def ctx():
    import cigwas_b200

    c = cigwas_b200.Context(0)
    yield c
    c.close()


def _case():
    import cigwas_b200
    from cigwas_b200 import synth

    Cm = synth.sem_correlation(220, 5, stream=17, num_samples=NS, causal_per_trait=8)
    n = Cm.shape[0]
    ti = np.zeros(n, np.int32)
    ti[220:] = 1
    return Cm, n, cigwas_b200.hetcor_threshold(ALPHA), ti


@pytest.mark.parametrize("case", ["g_absent", "g_ones", "late_zero_in_g", "late_value_in_n", "both"])
def test_speculative_start_and_its_fallback(ctx, oracle, case):
    Cm, n, th, ti = _case()
    N = np.full((n, n), float(NS), np.float32)
    G = np.ones((n, n), np.int32)
    if case in ("late_zero_in_g", "both"):
        G[n - 1, n - 2] = G[n - 2, n - 1] = 0  # a trait-trait pair that is not a candidate
        G[n - 3, 5] = G[5, n - 3] = 0
    if case in ("late_value_in_n", "both"):
        N[n - 1, n - 3] = N[n - 3, n - 1] = 40000.0
    for _ in range(2):  # twice: the helper thread and its team are reused
        rg = ctx.hetcor_skeleton(Cm, None if case == "g_absent" else G, N, th, 14, ti)
        ro = oracle.hetcor_skeleton(Cm, G, N, th, 14, ti, near_band=NEAR_BAND)
        assert rg["l"] == ro["l"]
        _assert_adj_equal(rg["G"], ro["G"], ro["near"])
        assert rg["stats"]["tests_canonical"][:15] == ro["stats"][:15].tolist()
    up = rg["stats"]["h2d_bytes"]
    mats = 1 + (case in ("late_zero_in_g", "both")) + (case in ("late_value_in_n", "both"))
    assert n * n * 4 * mats <= up < n * n * 4 * mats + 8 * n, "only the matrices that had to travel were uploaded"
