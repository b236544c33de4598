"""BASELINE.json configs[0] on the CPU side: the simulate_dag.R generator restated in cigwas_b200.synth.simulate_dag
(simulation/simulate_dag.R:3-124) is deterministic, and the oracle's skeleton of it (alpha 1e-4, max level 3, both
engines) behaves as a PC skeleton of that DAG should.  north_star names pcalg::skeleton as the CPU cross-check for
this config; R is not installed here or on the GPU box, so the oracle stands alone.  The GPU
side of the same case is tests/test_zzzz_c1_gpu.py."""
import numpy as np
import pytest

ALPHA, NS, SNP, TR = 1e-4, 10000, 1600, 10


@pytest.fixture(scope="module")
def c1():
    from cigwas_b200 import synth

    return synth.simulate_dag(n=NS, snp=SNP, tr=TR)


def test_generator_is_deterministic_and_shaped(c1):
    from cigwas_b200 import synth

    C, truth = c1
    assert C.shape == (SNP + TR, SNP + TR) and C.dtype == np.float32
    assert (np.diag(C) == 1).all() and np.array_equal(C, C.T) and np.abs(C).max() <= 1
    C2, truth2 = synth.simulate_dag(n=NS, snp=SNP, tr=TR)
    assert np.array_equal(C, C2) and np.array_equal(truth, truth2)
    # deg = 3: about 3 children per marker, plus the trait block (simulate_dag.R:15-16)
    assert 2000 < truth.sum() // 2 < 2900
    assert not np.array_equal(C, synth.simulate_dag(n=NS, snp=SNP, tr=TR, stream=2)[0])


def test_oracle_skeleton_of_c1(oracle, c1):
    C, truth = c1
    n = C.shape[0]
    th = oracle.hetcor_threshold(ALPHA)
    N = np.full((n, n), float(NS), np.float32)
    ti = np.zeros(n, np.int32)
    ti[SNP:] = 1  # cuskss: markers 0, traits 1 (cli.cpp:37-42)
    ones = np.ones((n, n), np.int32)
    prev = None
    for level in range(4):
        G = oracle.hetcor_skeleton(C, ones, N, th, level, ti)["G"].astype(bool)
        assert np.array_equal(G, G.T) and not G.diagonal().any()
        if prev is not None:
            assert not (G & ~prev).any()  # a level only removes edges
        prev = G
    tp, fp = (prev & truth).sum() // 2, (prev & ~truth).sum() // 2
    assert tp > 0.7 * (truth.sum() // 2) and fp < 0.06 * (prev.sum() // 2), (tp, fp)
    # the individual-level engine with the per-level thresholds of n = 10 000 keeps the same edges on this case
    r = oracle.skeleton(C, oracle.threshold_array(NS, ALPHA), 3)
    assert np.array_equal(r["G"].astype(bool), prev)
    # every removed edge above level 0 carries a separating set of its level's size, none of it an end of the edge
    sep = r["SepSet"]
    removed = np.argwhere((r["G"] == 0) & (sep[:, :, 0] >= 0))
    assert len(removed) > 100
    for x, y in removed[:500]:
        s = sep[x, y][sep[x, y] >= 0]
        assert 1 <= len(s) <= 3 and x not in s and y not in s
