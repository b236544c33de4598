"""Parity at the sizes BASELINE.json's configs name ("What's weak" 1): the smaller parity cases of
tests/test_skeleton_gpu.py / test_corr_gpu.py stop at 417 variables and n = 4 100 samples; these run the CUDA path
through the C ABI on the full shapes and compare with the CPU oracle.

  C2  cuskss single LD block, 3 000 markers + 17 traits (n_var 3 017), max level 14: both engines
  C2d the same shape from SURVEY 8d's generator (AR(1) LD inside sub-blocks of 50, 30 causal markers per trait), to
      max level 5: its trait rows keep ~30 neighbours, so levels 6..14 are 10^8..10^10 tests each -- hours for the
      oracle; the deep levels of that block are covered by the cross-kernel test below instead
  C5  cuskss-merged, 20 000 markers + 17 traits (n_var 20 017), heterologous N with NaN, time index
  C3  BED -> correlations at n = 458 000 samples (m = 2 048 / 1 024 markers, 17 traits with NA): >= 400 sampled
      3 x 3 tables bit-exact, every mxp / pxp value and the sampled mxm values within 1e-6, with and without
      missing calls

Bars as everywhere (north_star): integer tables bit-exact, correlations 1e-6 relative, adjacency identical outside
the 1e-9 band around the threshold, canonical test counts exact.
"""
import numpy as np
import pytest

from tests.test_skeleton_gpu import ALPHA, NEAR_BAND, NS, _assert_adj_equal

pytestmark = pytest.mark.gpu

M2, P2 = 3000, 17


@pytest.fixture(scope="module")
def ctx():
    import cigwas_b200

    c = cigwas_b200.Context(0)
    yield c
    c.close()


def _c2(dense):
    import bench
    from cigwas_b200 import synth

    if dense:
        return synth.c2_dense_block(M2, P2, stream=2, num_samples=NS)
    return synth.sem_correlation(M2, P2, stream=2, num_samples=NS, **bench.C2_KW)


@pytest.mark.parametrize("dense", [False, True])
def test_c2_full_size_summary_statistic_engine(ctx, oracle, dense):
    """the bench's own block (stream 2) and the SURVEY 8d generator, cuskss semantics (traits time 1)"""
    Cm = _c2(dense)
    n = Cm.shape[0]
    th = oracle.hetcor_threshold(ALPHA)
    N = np.full((n, n), float(NS), np.float32)
    ti = np.zeros(n, np.int32)
    ti[M2:] = 1
    ml = 5 if dense else 14
    ro = oracle.hetcor_skeleton(Cm, np.ones((n, n), np.int32), N, th, ml, ti, near_band=NEAR_BAND)
    rg = ctx.hetcor_skeleton(Cm, None, N, th, ml, ti)
    _assert_adj_equal(rg["G"], ro["G"], ro["near"])
    assert rg["l"] == ro["l"]
    assert rg["stats"]["tests_canonical"][: min(ro["l"], ml) + 1] == ro["stats"][: min(ro["l"], ml) + 1].tolist()
    assert (rg["G"] == rg["G"].T).all() and np.diag(rg["G"]).sum() == 0


def test_c2_full_size_individual_level_engine(ctx, oracle):
    Cm = _c2(False)
    Th = oracle.threshold_array(NS, ALPHA)
    ro = oracle.skeleton(Cm, Th, 14, near_band=NEAR_BAND)
    rg = ctx.skeleton(Cm, Th, 14, want_pmax=True, want_dense_sepset=True)
    ndiff = _assert_adj_equal(rg["G"], ro["G"], ro["near"])
    assert rg["l"] == ro["l"]
    assert rg["stats"]["tests_canonical"][: min(ro["l"], 14) + 1] == ro["stats"][: min(ro["l"], 14) + 1].tolist()
    if ndiff == 0 and len(ro["near"]) == 0:
        assert (rg["SepSet"] == ro["SepSet"]).all()
        assert np.allclose(rg["pMax"], ro["pMax"], rtol=1e-5, atol=1e-7)


def test_c5_merged_het_full_size(ctx, oracle):
    """BASELINE.json configs[4] on one GPU: 20 017 variables, per-entry effective sample sizes with NaN, traits at
    time 1.  The reference cannot run this shape (n > 12 288, cuPC-S.cu:140-144); the oracle can."""
    from cigwas_b200 import synth

    loci, per, p = 400, 50, 17
    m = loci * per
    n = m + p
    Cm = synth.merged_correlation(loci, per, p, stream=5, num_samples=NS)
    N = synth.het_sample_sizes(n, m, NS, stream=5, lo=50000, na_rate=0.005)
    assert np.isnan(N).any()
    ti = np.zeros(n, np.int32)
    ti[m:] = 1
    th = oracle.hetcor_threshold(ALPHA)
    ro = oracle.hetcor_skeleton(Cm, np.ones((n, n), np.int32), N, th, 14, ti, near_band=NEAR_BAND)
    rg = ctx.hetcor_skeleton(Cm, None, N, th, 14, ti)
    _assert_adj_equal(rg["G"], ro["G"], ro["near"])
    assert rg["l"] == ro["l"]
    assert rg["stats"]["tests_canonical"][: min(ro["l"], 14) + 1] == ro["stats"][: min(ro["l"], 14) + 1].tolist()


@pytest.mark.parametrize("m,miss", [(2048, 0.0), (1024, 0.002)])
def test_c3_size_correlations(ctx, oracle, m, miss):
    """n = 458 000 samples: the 458 112-byte-wide tensor map, int32 accumulators near their largest values and the
    limb fixed-point mxp path, all at the size whose throughput scripts/bench_corr.py quotes."""
    from cigwas_b200 import synth

    n, p = 458000, 17
    bed = synth.random_bed(n, m, stream=3, missing_rate=miss)
    rng = synth.rng_for(503)
    ph = rng.standard_normal((p, n)).astype(np.float32)
    for t in range(3):  # 1 % NA in three traits
        ph[t, rng.random(n) < 0.01] = np.nan
    means, stds, modes, counts = ctx.bed_marker_stats(bed, m, n)
    om, os_, omo, oc = oracle.bed_col_stats(bed, m, n)
    assert (counts == oc).all(), "genotype counts must be bit-exact"
    assert (means == om).all() and (stds == os_).all() and (modes == omo).all()
    r = ctx.corr_block(bed, ph, m, n, p, means, stds, want_tables=True)
    assert r["stats"]["parts"] == (3 if miss > 0 else 2)
    rs = np.random.default_rng(m)
    ia = rs.integers(0, m, size=600)
    ib = rs.integers(0, m, size=600)
    keep = ia != ib
    a_, b_ = np.minimum(ia, ib)[keep][:450], np.maximum(ia, ib)[keep][:450]
    assert len(a_) >= 400
    for a, b in zip(a_.tolist(), b_.tolist()):
        lin = a * (2 * m - a - 1) // 2 + (b - a - 1)
        t = oracle.pair_table(bed, n, a, b)
        assert (r["tables"][lin] == t).all(), (a, b)
        expect = oracle.npn_from_table(t)
        got = float(r["mxm"][lin])
        assert (np.isnan(expect) and np.isnan(got)) or abs(got - expect) <= 1e-6 * abs(expect) + 2e-7, (a, b, got, expect)
    _, omxp, opxp = oracle.corr_npn(bed, ph.reshape(-1), m, n, p, means, stds, want_mxm=False)
    x, y = np.asarray(r["mxp"], np.float64), np.asarray(omxp, np.float64)
    ok = np.isclose(x, y, rtol=1e-6, atol=2e-7) | (np.isnan(x) & np.isnan(y))
    assert ok.all(), f"mxp: max abs diff {np.nanmax(np.abs(x - y))}"
    assert np.allclose(r["pxp"], opxp, rtol=1e-6, atol=2e-7)
