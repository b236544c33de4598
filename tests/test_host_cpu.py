"""CPU tests of the C++ host layer (no device): the cuskss input assembly against the reference's own
test files and against the numpy restatement, variable selection against the reference's golden sets,
the phenotype loader, and the `mps` argv / exit-status contract on input errors."""
import base64
import ctypes as C
import os
import subprocess

import numpy as np
import pytest

from tests import host_ref

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
MPS = os.path.join(ROOT, "ci-gwas_b200", "mps")


@pytest.fixture(scope="module")
def lib():
    import cigwas_b200

    lib = cigwas_b200.load_library()
    fpp = C.POINTER(C.POINTER(C.c_float))
    lib.cusk_cuskss_square_inputs.argtypes = [C.c_char_p] * 6 + [C.c_int, C.c_char_p, C.c_float, C.POINTER(C.c_int),
                                              C.POINTER(C.c_int), fpp, fpp]
    lib.cusk_free.argtypes = [C.c_void_p]
    lib.cusk_subset_variables.argtypes = [C.POINTER(C.c_int), C.c_int, C.c_int, C.c_int, C.POINTER(C.c_int)]
    lib.cusk_load_phen.argtypes = [C.c_char_p, C.POINTER(C.c_int64), C.POINTER(C.c_int64), fpp]
    return lib


@pytest.fixture(scope="module")
def ref_files(golden, tmp_path_factory):
    d = tmp_path_factory.mktemp("ref_files")
    for name, b in golden["files_b64"].items():
        (d / name).write_bytes(base64.b64decode(b))
    return str(d)


def square(lib, mxm, mxp, mxp_se, pxp, pxp_se, blockfile, bix, ixs, ns):
    nv, p = C.c_int(0), C.c_int(0)
    Cp, Np = C.POINTER(C.c_float)(), C.POINTER(C.c_float)()
    enc = lambda s: s.encode() if s else b"NULL"
    rc = lib.cusk_cuskss_square_inputs(enc(mxm), enc(mxp), enc(mxp_se), enc(pxp), enc(pxp_se), enc(blockfile), bix,
                                       enc(ixs), C.c_float(ns), C.byref(nv), C.byref(p), C.byref(Cp), C.byref(Np))
    if rc != 0:
        raise RuntimeError(lib.cusk_last_error().decode())
    n = nv.value
    Cm = np.ctypeslib.as_array(Cp, (n, n)).copy()
    Nm = np.ctypeslib.as_array(Np, (n, n)).copy() if Np else None
    lib.cusk_free(Cp)
    if Np:
        lib.cusk_free(Np)
    return Cm, Nm, p.value


def test_square_inputs_reference_files(lib, ref_files):
    """cusk/tests/cuskss_tests.cpp:85-209: merged (marker indices) and block selection give the same 6x6."""
    f = lambda n: os.path.join(ref_files, n)
    Cm, Nm, p = square(lib, f("small_mxm.bin"), f("marker_trait_summary_stats.txt"), None, f("trait_summary_stats.txt"),
                       None, None, 0, f("marker_indices.bin"), 500000.0)
    Cb, Nb, _ = square(lib, f("small_mxm.bin"), f("marker_trait_summary_stats.txt"), None, f("trait_summary_stats.txt"),
                       None, f("blocks.txt"), 0, None, 500000.0)
    assert p == 3 and Cm.shape == (6, 6) and Nm is None and Nb is None
    assert (Cm == Cb).all()
    mxm = host_ref.load_mxm(f("small_mxm.bin"))
    mxp, _ = host_ref.load_mxp(f("marker_trait_summary_stats.txt"), None, [0, 1, 2])
    pxp, _ = host_ref.load_pxp(f("trait_summary_stats.txt"), None, 500000.0)
    Ce, _ = host_ref.square_inputs(mxm, mxp, None, pxp, None, 500000.0, False)
    assert (Cm == Ce).all()
    assert (Cm == Cm.T).all()
    np.testing.assert_allclose(Cm[2, 3:], [0.0005, 0.0001, -0.01], rtol=1e-6)
    np.testing.assert_allclose(Cm[3, 4], 0.0608594558771734, rtol=1e-6)


def test_square_inputs_het_block_with_offset_and_na(lib, tmp_path):
    """hetcor inputs: SE -> N = ((1 - r^2)/se)^2 (marker_trait_summary_stats.cpp:162-164), NA -> r = 0, N = NaN;
    the block's rows are selected by its global offset in the mxp file (marker_block.h:52-56)."""
    from cigwas_b200 import synth

    m, p = 40, 4
    Cm = synth.sem_correlation(m, p, stream=71, num_samples=50000, causal_per_trait=6)
    N = synth.het_sample_sizes(m + p, m, 50000, stream=72, lo=5000, na_rate=0.0)
    d = str(tmp_path)
    host_ref.write_cuskss_inputs(d, Cm, m, p, het_N=N, na_frac=0.1, rng=np.random.default_rng(3), extra_rows_before=25)
    with open(os.path.join(d, "blocks.txt"), "w") as f:
        f.write("0\t0\t24\n1\t0\t%d\n" % (m - 1))
    f = lambda n: os.path.join(d, n)
    Cg, Ng, pp = square(lib, f("mxm.bin"), f("mxp.txt"), f("mxp_se.txt"), f("pxp.txt"), f("pxp_se.txt"), f("blocks.txt"), 1,
                        None, 50000.0)
    assert pp == p and Ng is not None
    mxm = host_ref.load_mxm(f("mxm.bin"))
    mxp, mess = host_ref.load_mxp(f("mxp.txt"), f("mxp_se.txt"), list(range(25, 25 + m)))
    pxp, pess = host_ref.load_pxp(f("pxp.txt"), f("pxp_se.txt"), 50000.0)
    Ce, Ne = host_ref.square_inputs(mxm, mxp, mess, pxp, pess, 50000.0, True)
    assert (Cg == Ce).all()
    assert np.array_equal(Ng, Ne, equal_nan=True)
    assert np.isnan(Ng).any() and (Cg[np.isnan(Ng)] == 0).all()
    ok = ~np.isnan(Ng[:m, m:])
    np.testing.assert_allclose(Ng[:m, m:][ok], N[:m, m:][ok], rtol=2e-3)  # text round trip of se
    # marker-indices selection of the same rows gives the same matrices
    np.arange(25, 25 + m, dtype=np.int32).tofile(f("ixs.bin"))
    Cg2, Ng2, _ = square(lib, f("mxm.bin"), f("mxp.txt"), f("mxp_se.txt"), f("pxp.txt"), f("pxp_se.txt"), None, 0,
                         f("ixs.bin"), 50000.0)
    assert (Cg2 == Cg).all() and np.array_equal(Ng2, Ng, equal_nan=True)


def test_square_inputs_errors(lib, ref_files, tmp_path):
    f = lambda n: os.path.join(ref_files, n)
    np.zeros(10, np.float32).tofile(str(tmp_path / "mxm4.bin"))  # 4 markers vs 3 mxp rows
    with pytest.raises(RuntimeError, match="markers seem to differ"):
        square(lib, str(tmp_path / "mxm4.bin"), f("marker_trait_summary_stats.txt"), None, f("trait_summary_stats.txt"),
               None, f("blocks.txt"), 0, None, 1000.0)
    with pytest.raises(RuntimeError, match="not found"):
        square(lib, f("nope.bin"), f("marker_trait_summary_stats.txt"), None, f("trait_summary_stats.txt"), None,
               f("blocks.txt"), 0, None, 1000.0)


def test_subset_variables_golden(lib, golden):
    """cusk/tests/parents_tests.cpp:5-39 on the 20-node graph of parent_set_test_set.h"""
    ps = golden["parent_set"]
    nm, npn = ps["num_markers"], ps["num_phen"]
    nv = nm + npn
    G = np.array(ps["adj"], np.int32).reshape(nv, nv)
    for depth, key in [(0, "d0"), (1, "d1"), (2, "d2")]:
        out = np.zeros(nv, np.int32)
        assert lib.cusk_subset_variables(G.ctypes.data_as(C.POINTER(C.c_int)), nv, nm, depth,
                                         out.ctypes.data_as(C.POINTER(C.c_int))) == 0
        assert sorted(np.nonzero(out)[0].tolist()) == sorted(ps[key]), depth
        assert sorted(np.nonzero(out)[0].tolist()) == host_ref.subset_variables(G, nm, depth)


def test_subset_variables_random(lib):
    rng = np.random.default_rng(5)
    for nm, npn, dens in [(60, 4, 0.03), (200, 7, 0.01), (30, 1, 0.2)]:
        nv = nm + npn
        G = (rng.random((nv, nv)) < dens).astype(np.int32)
        G = np.triu(G, 1)
        G = G + G.T
        for depth in range(0, 4):
            out = np.zeros(nv, np.int32)
            lib.cusk_subset_variables(G.ctypes.data_as(C.POINTER(C.c_int)), nv, nm, depth,
                                      out.ctypes.data_as(C.POINTER(C.c_int)))
            assert np.nonzero(out)[0].tolist() == host_ref.subset_variables(G, nm, depth)


def test_load_phen(lib, ref_files):
    """src/phen.cpp:27-74: header skipped, two id columns dropped, NA -> NaN, column-major"""
    path = os.path.join(ref_files, "with_nan.phen")
    ns, npn = C.c_int64(0), C.c_int64(0)
    dp = C.POINTER(C.c_float)()
    assert lib.cusk_load_phen(path.encode(), C.byref(ns), C.byref(npn), C.byref(dp)) == 0
    got = np.ctypeslib.as_array(dp, (npn.value, ns.value)).copy()
    lib.cusk_free(dp)
    rows = [ln.split()[2:] for ln in open(path).read().split("\n")[1:] if ln.split()]
    exp = np.array([[np.nan if v == "NA" else np.float32(float(v)) for v in r] for r in rows], np.float32).T
    assert got.shape == exp.shape and np.array_equal(got, exp, equal_nan=True)
    assert np.isnan(got).any()


def test_mps_exit_status_contract(ref_files, tmp_path):
    """apps/mps.cpp:17-30 (usage, status 0), io.cpp:399-406 (missing path: message + status 1)"""
    assert os.path.exists(MPS), "run __graft_entry__.build() first"
    r = subprocess.run([MPS], capture_output=True, text=True)
    assert r.returncode == 0 and "usage: mps <command>" in r.stdout
    r = subprocess.run([MPS, "prep", str(tmp_path / "nothing")], capture_output=True, text=True)
    assert r.returncode == 1 and "file or directory not found" in r.stdout
    f = lambda n: os.path.join(ref_files, n)
    args = [MPS, "cuskss", f("small_mxm.bin"), f("marker_trait_summary_stats.txt"), "NULL", f("missing_pxp.txt"), "NULL",
            "NULL", "0", f("blocks.txt"), "NULL", "0.0001", "3", "1", "1", "500000", str(tmp_path)]
    r = subprocess.run(args, capture_output=True, text=True)
    assert r.returncode == 1 and "missing_pxp.txt" in r.stdout
    r = subprocess.run([MPS, "cusk", "a", "b"], capture_output=True, text=True)
    assert r.returncode == 1 and "usage: mps cusk" in r.stdout
