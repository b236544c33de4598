"""The CPU oracle against every golden vector the reference's own tests hold for the path
(SURVEY 8c).  CPU only."""
import base64
import math

import numpy as np


def test_threshold_golden(oracle, golden):
    t = golden["threshold"]
    th = oracle.threshold_array(t["n"], t["alpha"])
    assert abs(th[0] - t["th0"]) < t["tol"]
    # cross-check qnorm against scipy (same maths as cusk_postprocessing/sepselect.py:25-26)
    from scipy.stats import norm

    for p in [5e-5, 5e-9, 0.025, 0.3, 0.5, 0.9, 1 - 1e-7]:
        assert abs(oracle.qnorm(p) - norm.ppf(p)) <= 1e-12 * max(1.0, abs(norm.ppf(p)))
    assert abs(oracle.hetcor_threshold(1e-4) - 3.8906) < 1e-3


def test_skeleton_c_n10_both_engines(oracle, golden):
    g = golden["cupc_n10"]
    n = g["n"]
    Cm = np.array(g["C"], np.float32).reshape(n, n)
    A = np.array(g["A"], np.int32).reshape(n, n)
    Th = oracle.threshold_array(g["sample_size"], g["alpha"])
    r = oracle.skeleton(Cm, Th, 14)
    assert (r["G"] == A).all()
    th = oracle.hetcor_threshold(g["alpha"])
    N = np.full((n, n), float(g["sample_size"]), np.float32)
    r2 = oracle.hetcor_skeleton(Cm, np.ones((n, n), np.int32), N, th, 14, np.zeros(n, np.int32))
    assert (r2["G"] == A).all()
    # sepsets: every removed ordered pair with a recorded set has a set drawn from other nodes
    S = r["SepSet"]
    for i in range(n):
        for j in range(n):
            s = S[i, j][S[i, j] >= 0]
            assert i not in s and j not in s


def test_corr_small_golden(oracle, golden):
    b = golden["bmt"]
    bed = np.array(b["bed"], np.uint8)
    phen = np.array(b["phen"], np.float32)
    mxm, mxp, pxp = oracle.corr_npn(bed, phen, b["m"], b["n"], b["p"], b["mean"], b["std"])
    assert np.allclose(mxp, b["mxp_expected"], atol=b["tol"])
    assert np.allclose(pxp, b["pxp_expected"], atol=b["tol"])
    assert np.allclose(mxm, b["mxm_npn_expected"], atol=b["tol"])


def test_corr_bmt2_golden(oracle, golden):
    b = golden["bmt2"]
    bed = np.array(b["bed"], np.uint8)
    phen = np.array(b["phen"], np.float32)
    mxm, mxp, pxp = oracle.corr_npn(bed, phen, b["m"], b["n"], b["p"], b["mean"], b["std"])
    assert np.allclose(mxm, b["mxm_npn_expected"], atol=b["tol"])
    assert np.allclose(mxp, b["mxp_expected"], atol=b["tol"])
    assert np.allclose(pxp, b["pxp_expected"], atol=b["tol"])


def test_prep_small_bed(oracle, golden):
    raw = base64.b64decode(golden["files_b64"]["small.bed"])
    assert raw[:3] == bytes([0x6C, 0x1B, 0x01])
    p = golden["prep_small"]
    bed = np.frombuffer(raw[3:], np.uint8)
    means, stds, modes, counts = oracle.bed_col_stats(bed, p["m"], p["n"])
    assert np.allclose(means, p["means"], atol=p["tol"])
    assert np.allclose(stds, p["stds"], atol=p["tol"])
    assert (counts.sum(axis=1) == p["n"]).all()


def test_with_nan_pxp(oracle, golden):
    txt = base64.b64decode(golden["files_b64"]["with_nan.phen"]).decode()
    lines = txt.strip().split("\n")[1:]
    rows = [[float("nan") if f in ("NA", "nan", "NaN") else float(f) for f in ln.split()[2:]] for ln in lines]
    ph = np.array(rows, np.float32)  # samples x traits
    n, p = ph.shape
    phen = np.ascontiguousarray(ph.T).reshape(-1)
    bed = np.zeros(((n + 3) // 4), np.uint8)
    _, _, pxp = oracle.corr_npn(bed, phen, 1, n, p, [0.0], [1.0], want_mxm=False)
    assert np.allclose(pxp, golden["with_nan_pxp_expected"], atol=1e-5)


def test_parent_set_golden(oracle, golden):
    ps = golden["parent_set"]
    nv = ps["num_markers"] + ps["num_phen"]
    G = np.array(ps["adj"], np.int32).reshape(nv, nv)
    for depth, key in [(0, "d0"), (1, "d1"), (2, "d2")]:
        sel = oracle.subset_variables(G, nv, ps["num_markers"], depth)
        assert list(sel) == ps[key]


def test_inverse_restatements_agree(oracle):
    rng = np.random.default_rng(5)
    for m in [2, 3, 4, 7, 14]:
        X = rng.standard_normal((200, m))
        A = np.corrcoef(X.T)
        ref = np.linalg.inv(A)
        assert np.allclose(oracle.inv_svd(A), ref, rtol=1e-9, atol=1e-9)
        if m <= 3:
            assert np.allclose(oracle.pinv_courrieu(A), ref, rtol=1e-9, atol=1e-9)
    # NOTE: with the reference's tolerance (1e-20 * min diag) a numerically singular M2 is NOT
    # reliably detected (the rounding residue of the rejected pivot is ~1e-16 >> tol), so singular
    # conditioning sets are outside the parity domain at every level (DESIGN.md "Parity domain").


def test_combinations(oracle):
    import itertools

    for N, P in [(5, 2), (7, 3), (9, 4), (6, 6)]:
        combs = list(itertools.combinations(range(1, N + 1), P))
        assert oracle.binom(N, P) == len(combs)
        for L, c in enumerate(combs, start=1):
            assert tuple(oracle.ith_combination(N, P, L)) == c
    assert oracle.binom(60, 14) == math.comb(60, 14)  # overflows int32 in the reference


def test_skeleton_matches_bruteforce_numpy(oracle):
    """Independent brute force in numpy (np.linalg.inv) on a small random DAG-like matrix."""
    import itertools

    rng = np.random.default_rng(11)
    n, ns = 12, 2000
    B = np.triu(rng.uniform(-0.6, 0.6, (n, n)) * (rng.random((n, n)) < 0.3), 1)
    X = rng.standard_normal((ns, n)) @ np.linalg.inv(np.eye(n) - B)
    Cm = np.corrcoef(X.T).astype(np.float32)
    Th = oracle.threshold_array(ns, 0.01)
    r = oracle.skeleton(Cm, Th, 14)
    Cd = Cm.astype(np.float64)
    np.fill_diagonal(Cd, 1.0)
    G = (np.abs(np.arctanh(np.clip(Cd - np.eye(n), -0.999999999, 0.999999999))) >= Th[0]).astype(int)
    np.fill_diagonal(G, 0)
    l = 1
    while True:
        adj = [list(np.nonzero(G[i])[0]) for i in range(n)]
        if max(len(a) for a in adj) - 1 < l or l > 14:
            break
        rem = []
        for x in range(n):
            for y in adj[x]:
                others = [v for v in adj[x] if v != y]
                for S in itertools.combinations(others, l):
                    idx = [x, y] + list(S)
                    P = np.linalg.inv(Cd[np.ix_(idx, idx)])
                    rho = -P[0, 1] / math.sqrt(P[0, 0] * P[1, 1])
                    if abs(math.atanh(rho)) < Th[l]:
                        rem.append((x, y))
                        break
        for x, y in rem:
            G[x, y] = G[y, x] = 0
        l += 1
    assert (r["G"] == G).all()


def test_ref_quirk_counters(oracle):
    """orc_ref_quirks on constructed rows: the reference's NoEdgeFlag early exit (hetcor-cuPC-S.cu:550-589) needs a
    whole 64-rank pass of one CTA without a (Y, S) that the time rule allows while later passes still hold one; with
    markers at time 0 and traits at time 1 that cannot happen (any trait neighbour outside S makes S admissible), with
    the FIRST trait as the latest variable every set that contains it is inadmissible and the row is cut short from
    level 3 on; int32 ranks overflow at C(40, 14)."""
    n, m = 60, 43
    G = np.zeros((n, n), np.int32)
    G[0, 1:] = 1
    G[1:, 0] = 1
    ti = np.zeros(n, np.int32)
    ti[m:] = 1
    for l in (2, 3, 4):
        q = oracle.ref_quirks(G, ti, l)
        assert q["rows_cut_short_by_noedgeflag"] == 0 and q["rows_int32_rank_overflow"] == 0
    ti[m] = 3
    assert oracle.ref_quirks(G, ti, 2)["rows_cut_short_by_noedgeflag"] == 0
    q = oracle.ref_quirks(G, ti, 3)
    assert q["rows_cut_short_by_noedgeflag"] == 1 and q["sets_never_evaluated"] == 256
    G2 = np.zeros((45, 45), np.int32)
    G2[0, 1:41] = 1
    G2[1:41, 0] = 1
    q = oracle.ref_quirks(G2, np.zeros(45, np.int32), 14, cap=1000)
    assert q["rows_int32_rank_overflow"] == 1 and q["rows_not_analysed"] == 1
