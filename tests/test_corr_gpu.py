"""Parity of the CUDA correlation front end (through the C ABI) against the reference's golden vectors
and the CPU oracle.  Bars (BASELINE.json north_star): genotype counts and integer dot products
(contingency tables) bit-exact; correlations within 1e-6 relative of the FP64 oracle."""
import base64
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

RTOL = 1e-6


@pytest.fixture(scope="module")
def ctx():
    import cigwas_b200

    c = cigwas_b200.Context(0)
    yield c
    c.close()


def _close(a, b, rtol=RTOL, atol=2e-7):
    # outputs are float32 (the reference's output type): 1e-6 relative + half a float ulp of O(1) values
    a = np.asarray(a, np.float64)
    b = np.asarray(b, np.float64)
    ok = np.isclose(a, b, rtol=rtol, atol=atol) | (np.isnan(a) & np.isnan(b))
    assert ok.all(), f"max abs diff {np.nanmax(np.abs(a - b))} at {np.argwhere(~ok)[:5].tolist()}"


def test_golden_small_sets(ctx, golden):
    """cusk/tests/corr_tests.cpp:211-289 (10 x 3 x 2) and :141-185,291-336 (100 x 7 x 5), tolerance 1e-5."""
    for key in ("bmt", "bmt2"):
        b = golden[key]
        bed = np.array(b["bed"], np.uint8)
        phen = np.array(b["phen"], np.float32)
        r = ctx.corr_block(bed, phen, b["m"], b["n"], b["p"], b["mean"], b["std"])
        assert np.allclose(r["mxm"], b["mxm_npn_expected"], atol=b["tol"])
        assert np.allclose(r["mxp"], b["mxp_expected"], atol=b["tol"])
        assert np.allclose(r["pxp"], b["pxp_expected"], atol=b["tol"])
        if "mxm_pearson_expected" in b:
            r2 = ctx.corr_block(bed, phen, b["m"], b["n"], b["p"], b["mean"], b["std"], mode=1)
            assert np.allclose(r2["mxm"], b["mxm_pearson_expected"], atol=b["tol"])


def test_golden_prep_small_bed(ctx, oracle, golden):
    """cusk/tests/prep_tests.cpp:44-73"""
    raw = base64.b64decode(golden["files_b64"]["small.bed"])
    p = golden["prep_small"]
    bed = np.frombuffer(raw[3:], np.uint8)
    means, stds, modes, counts = ctx.bed_marker_stats(bed, p["m"], p["n"])
    assert np.allclose(means, p["means"], atol=p["tol"])
    assert np.allclose(stds, p["stds"], atol=p["tol"])
    om, os_, omo, oc = oracle.bed_col_stats(bed, p["m"], p["n"])
    assert (means == om).all() and (stds == os_).all() and (modes == omo).all() and (counts == oc).all()


def test_with_nan_phen(ctx, golden):
    """cusk/tests/corr_tests.cpp:187-209"""
    txt = base64.b64decode(golden["files_b64"]["with_nan.phen"]).decode()
    lines = txt.strip().split("\n")[1:]
    rows = [[float("nan") if f in ("NA", "nan", "NaN") else float(f) for f in ln.split()[2:]] for ln in lines]
    ph = np.array(rows, np.float32)
    n, p = ph.shape
    phen = np.ascontiguousarray(ph.T).reshape(-1)
    bed = np.zeros((n + 3) // 4, np.uint8)
    r = ctx.corr_block(bed, phen, 1, n, p, [0.0], [1.0], want_mxm=False)
    assert np.allclose(r["pxp"], golden["with_nan_pxp_expected"], atol=1e-5)


@pytest.mark.parametrize("n,m,p,miss,stream", [(1000, 70, 4, 0.0, 61), (777, 150, 5, 0.01, 62), (4100, 300, 3, 0.0, 63),
                                                (130, 45, 2, 0.05, 64), (515, 129, 0, 0.0, 65),
                                                # ceil(n / 4) % 4 == 0: 4-byte aligned columns (32-bit decode loads, the
                                                # shared-memory staged statistics kernel), with and without a ragged tail
                                                (4093, 97, 3, 0.01, 66), (1024, 64, 2, 0.0, 67)])
def test_random_block_vs_oracle(ctx, oracle, n, m, p, miss, stream):
    from cigwas_b200 import synth

    bed = synth.random_bed(n, m, stream=stream, missing_rate=miss)
    rng = synth.rng_for(stream + 500)
    ph = rng.standard_normal((p, n)).astype(np.float32)
    if p:
        ph[0, rng.random(n) < 0.02] = np.nan
    # statistics: exact
    means, stds, modes, counts = ctx.bed_marker_stats(bed, m, n)
    om, os_, omo, oc = oracle.bed_col_stats(bed, m, n)
    assert (counts == oc).all(), "genotype counts must be bit-exact"
    assert (means == om).all() and (stds == os_).all() and (modes == omo).all()
    # keep monomorphic markers out of the correlation comparison (0/0 in both implementations)
    r = ctx.corr_block(bed, ph, m, n, p, means, stds, want_tables=True)
    assert r["stats"]["parts"] == (3 if miss > 0 else 2)
    # integer dot products: every contingency table bit-exact
    lin = 0
    rs = np.random.default_rng(stream)
    pairs = [(a, b) for a in range(m) for b in range(a + 1, m)]
    sel = rs.choice(len(pairs), size=min(400, len(pairs)), replace=False)
    for k in sel:
        a, b = pairs[k]
        lin = a * (2 * m - a - 1) // 2 + (b - a - 1)
        assert (r["tables"][lin] == oracle.pair_table(bed, n, a, b)).all(), (a, b)
    assert (r["tables"].sum(axis=1) <= n).all()
    omxm, omxp, opxp = oracle.corr_npn(bed, ph.reshape(-1) if p else np.zeros(1, np.float32), m, n, p, means, stds)
    _close(r["mxm"], omxm)
    if p:
        _close(r["mxp"], omxp)
    if p > 1:
        _close(r["pxp"], opxp)
    # Pearson mode from the same tables
    r2 = ctx.corr_block(bed, ph, m, n, p, means, stds, mode=1)
    t = r["tables"].astype(np.float64)
    sab = t[:, 4] + 2 * (t[:, 5] + t[:, 7]) + 4 * t[:, 8]
    ia, ib = np.array(pairs).T
    mu, sd = means.astype(np.float64), stds.astype(np.float64)
    with np.errstate(divide="ignore", invalid="ignore"):
        expect = (sab / t.sum(axis=1) - mu[ia] * mu[ib]) / (sd[ia] * sd[ib])
    _close(r2["mxm"], expect)


def test_reference_named_entry_points(golden):
    """cu_corr_pearson_npn / cu_marker_phen_corr_pearson with the reference's argument order."""
    import cigwas_b200

    lib = cigwas_b200.load_library()
    b = golden["bmt2"]
    m, n, p = b["m"], b["n"], b["p"]
    bed = np.array(b["bed"], np.uint8)
    phen = np.array(b["phen"], np.float32)
    mean = np.array(b["mean"], np.float32)
    std = np.array(b["std"], np.float32)
    mxm = np.zeros(m * (m - 1) // 2, np.float32)
    mxp = np.zeros(m * p, np.float32)
    pxp = np.zeros(p * (p - 1) // 2, np.float32)
    ub, fp = C.POINTER(C.c_ubyte), C.POINTER(C.c_float)
    lib.cu_corr_pearson_npn.argtypes = [ub, fp, C.c_size_t, C.c_size_t, C.c_size_t, fp, fp, fp, fp, fp]
    lib.cu_corr_pearson_npn(bed.ctypes.data_as(ub), phen.ctypes.data_as(fp), m, n, p, mean.ctypes.data_as(fp),
                            std.ctypes.data_as(fp), mxm.ctypes.data_as(fp), mxp.ctypes.data_as(fp),
                            pxp.ctypes.data_as(fp))
    assert np.allclose(mxm, b["mxm_npn_expected"], atol=b["tol"])
    assert np.allclose(mxp, b["mxp_expected"], atol=b["tol"])
    assert np.allclose(pxp, b["pxp_expected"], atol=b["tol"])
    mxp2 = np.zeros(m * p, np.float32)
    lib.cu_marker_phen_corr_pearson.argtypes = [ub, fp, C.c_size_t, C.c_size_t, C.c_size_t, fp, fp, fp]
    lib.cu_marker_phen_corr_pearson(bed.ctypes.data_as(ub), phen.ctypes.data_as(fp), m, n, p,
                                    mean.ctypes.data_as(fp), std.ctypes.data_as(fp), mxp2.ctypes.data_as(fp))
    assert (mxp2 == mxp).all()


def test_corr_to_skeleton_device_pipeline(ctx, oracle):
    """BED block -> dense correlation matrix on the device -> Skeleton, nothing through the host."""
    import torch
    from cigwas_b200 import synth

    n, m, p = 2000, 120, 3
    bed = synth.random_bed(n, m, stream=71)
    rng = synth.rng_for(571)
    ph = rng.standard_normal((p, n)).astype(np.float32)
    means, stds, _, _ = ctx.bed_marker_stats(bed, m, n)
    nv = m + p
    d_bed = torch.from_numpy(bed).cuda()
    d_ph = torch.from_numpy(ph).cuda()
    d_mean = torch.from_numpy(means).cuda()
    d_std = torch.from_numpy(stds).cuda()
    d_C = torch.zeros((nv, nv), dtype=torch.float32, device="cuda")
    torch.cuda.synchronize()
    ctx.corr_block_dev(d_bed.data_ptr(), d_ph.data_ptr(), m, n, p, d_mean.data_ptr(), d_std.data_ptr(),
                       d_C.data_ptr())
    Cg = d_C.cpu().numpy()
    omxm, omxp, opxp = oracle.corr_npn(bed, ph.reshape(-1), m, n, p, means, stds)
    Co = np.eye(nv)
    iu = np.triu_indices(m, 1)
    Co[:m, :m][iu] = omxm
    Co[:m, m:] = omxp.reshape(m, p)
    Co[m:, m:][np.triu_indices(p, 1)] = opxp
    Co = np.triu(Co) + np.triu(Co, 1).T
    _close(Cg, Co)
    assert (Cg == Cg.T).all()


@pytest.mark.parametrize("n,m,miss", [(61, 40, 0.1), (2048, 33, 0.0), (2049, 31, 0.02), (8177, 100, 0.3), (16, 1, 0.0),
                                      (3, 5, 0.0), (6000, 257, 1.0)])
def test_marker_stats_shapes(ctx, oracle, n, m, miss):
    """prep-bed statistics on shapes that take each path of the round-2 kernels: 16-byte loads with a byte-wise head and
    tail in the counting kernel (columns at any alignment), the staged statistics kernel (ceil(n/4) % 4 == 0; several
    tiles, ragged last word, fewer than 32 markers in the last warp, all calls missing) and the byte-wise one (the rest):
    counts, means, standard deviations (the reference's sequential float sum) and modes equal the oracle's bit for bit"""
    from cigwas_b200 import synth

    bed = synth.random_bed(n, m, stream=300 + n % 97, missing_rate=min(miss, 0.99))
    if miss >= 1.0:
        bed = np.full_like(bed, 0x55)  # every call missing: mean = 0 / 0
    means, stds, modes, counts = ctx.bed_marker_stats(bed, m, n)
    om, os_, omo, oc = oracle.bed_col_stats(bed, m, n)
    assert (counts == oc).all()
    assert ((means == om) | (np.isnan(means) & np.isnan(om))).all()
    assert ((stds == os_) | (np.isnan(stds) & np.isnan(os_))).all()
    assert (modes == omo).all()


@pytest.mark.parametrize("n,m,miss,stream", [(4100, 1700, 0.0, 71), (2100, 1100, 0.004, 72)])
def test_gram_cta_pair_form_equals_one_cta_form(ctx, oracle, n, m, miss, stream):
    """blocks large enough for the K-panel meeting run the Gram kernel on CTA pairs (tcgen05.mma.cta_group::2, each CTA
    holding its A group and half of the B block; an odd number of 128-row groups, so the last pair's second CTA works on
    zero rows): every contingency table and every correlation equals the one-CTA-per-tile form's, and sampled tables
    the oracle's"""
    import os

    from cigwas_b200 import synth

    bed = synth.random_bed(n, m, stream=stream, missing_rate=miss)
    means, stds, modes, counts = ctx.bed_marker_stats(bed, m, n)
    ph = synth.rng_for(stream + 500).standard_normal((3, n)).astype(np.float32)
    saved = os.environ.get("CUSK_GRAM_PAIR")
    try:
        os.environ["CUSK_GRAM_PAIR"] = "0"
        r0 = ctx.corr_block(bed, ph, m, n, 3, means, stds, want_tables=True)
        os.environ["CUSK_GRAM_PAIR"] = "1"
        r1 = ctx.corr_block(bed, ph, m, n, 3, means, stds, want_tables=True)
    finally:
        if saved is None:
            os.environ.pop("CUSK_GRAM_PAIR", None)
        else:
            os.environ["CUSK_GRAM_PAIR"] = saved
    assert r1["stats"]["parts"] == (3 if miss > 0 else 2)
    assert (r1["tables"] == r0["tables"]).all()
    assert np.array_equal(r1["mxm"], r0["mxm"], equal_nan=True)
    assert np.array_equal(r1["mxp"], r0["mxp"], equal_nan=True)
    rs = np.random.default_rng(stream)
    for _ in range(200):
        a, b = sorted(rs.choice(m, size=2, replace=False).tolist())
        lin = a * (2 * m - a - 1) // 2 + (b - a - 1)
        assert (r1["tables"][lin] == oracle.pair_table(bed, n, a, b)).all(), (a, b)
