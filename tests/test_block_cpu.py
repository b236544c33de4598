"""`mps block` without a GPU: the oracle against the reference's golden vectors
(tests/golden/block_golden.json, made by tests/golden/make_block_golden.py from cusk/tests/blocking_tests.cpp,
cusk/tests/corr_tests.cpp:62-75 and their data headers), the product's host-side blocking (blocking.cpp, through the
C ABI: no device involved) against both, and the bit-plane routines of the banded kernel, compiled for the host,
against the oracle's contingency tables."""
import ctypes as C
import json
import os
import subprocess

import numpy as np
import pytest

from tests import oracle_lib

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def gold():
    with open(os.path.join(ROOT, "tests", "golden", "block_golden.json")) as f:
        g = json.load(f)
    with open(os.path.join(ROOT, "tests", "golden", "golden.json")) as f:
        g["bmt2"] = json.load(f)["bmt2"]
    return g


@pytest.fixture(scope="module")
def orc():
    return oracle_lib.load()


@pytest.fixture(scope="module")
def lib():
    from cigwas_b200 import lib as cl

    cl.load_library()
    return cl


def random_bed(rng, m, n, miss_rate=0.0, ld=0.9, miss_markers=None):
    """packed SNP-major columns with LD between neighbours; PLINK pads the last byte with 00 (= genotype 2)"""
    g = np.zeros((m, n), np.int8)
    maf = rng.uniform(0.05, 0.5, m)
    h = rng.random((2, n)) < maf[0]
    for i in range(m):
        if i:
            keep = rng.random((2, n)) < ld
            h = np.where(keep, h, rng.random((2, n)) < maf[i])
        g[i] = h.sum(0)
    code = np.array([3, 2, 0], np.uint8)[g]  # genotype 0 -> 11, 1 -> 10, 2 -> 00
    if miss_rate > 0:
        miss = rng.random((m, n)) < miss_rate
        if miss_markers is not None:
            miss[[i for i in range(m) if i not in miss_markers]] = False
        code[miss] = 1
    pad = (-n) % 4
    code = np.concatenate([code, np.zeros((m, pad), np.uint8)], 1).reshape(m, -1, 4)
    return (code[:, :, 0] | (code[:, :, 1] << 2) | (code[:, :, 2] << 4) | (code[:, :, 3] << 6)).astype(np.uint8)


# ---- oracle against the reference's golden vectors ----------------------------------------------------------

def test_oracle_banded_corr_golden(orc, gold):
    b = gold["bmt2"]
    exp = gold["bmt2_sparse_mcorrs"]
    got = orc.mcorrk_banded(np.array(b["bed"], np.uint8), b["m"], b["n"], exp["corr_width"])
    assert np.abs(got.ravel() - np.array(exp["expected"])).max() < exp["tol"]  # 1e-5, corr_tests.cpp:73


def test_oracle_hanning_smoothing_golden(orc, gold):
    v = np.array(gold["test_v"], np.float32)[: gold["smooth_n"]]
    got = orc.hanning_smoothing(v, gold["smooth_window"])
    assert np.abs(got - np.array(gold["test_v_smooth"])).max() < gold["smooth_tol"]  # 0.01, blocking_tests.cpp:52


def test_oracle_block_chr_golden(orc, gold):
    bc = gold["block_chr"]
    blocks, _ = orc.block_chr(np.array(gold["test_v"], np.float32), bc["max_block_size"])
    assert blocks.tolist() == bc["blocks"]  # exact, blocking_tests.cpp:9-38


# ---- product host side (blocking.cpp through the C ABI) -----------------------------------------------------

def test_host_hanning_smoothing(lib, orc, gold):
    v = np.array(gold["test_v"], np.float32)[: gold["smooth_n"]]
    got = lib.hanning_smoothing(v, gold["smooth_window"])
    assert np.abs(got - np.array(gold["test_v_smooth"])).max() < gold["smooth_tol"]
    assert np.array_equal(got, orc.hanning_smoothing(v, gold["smooth_window"]))
    # margins stay 0; window 1 is 0/0 in the reference's weight -> NaN everywhere, reproduced
    assert (got[:50] == 0).all() and (got[-50:] == 0).all()
    assert np.isnan(lib.hanning_smoothing(v[:10], 1)).all()


def test_host_block_chr_golden(lib, gold):
    bc = gold["block_chr"]
    blocks, window = lib.block_chr(np.array(gold["test_v"], np.float32), bc["max_block_size"])
    assert blocks.tolist() == bc["blocks"]
    assert window % 2 == 1


@pytest.mark.parametrize("m,max_block", [(1, 10), (2, 10), (3, 10), (7, 3), (50, 10), (997, 120), (3000, 400), (3000, 5000)])
def test_host_block_chr_matches_oracle(lib, orc, m, max_block):
    rng = np.random.default_rng(20261104 + m)
    # row sums with LD structure: smooth bumps + noise
    x = np.arange(m)
    v = (5 + 3 * np.sin(x / 17.0) + 2 * np.sin(x / 5.3 + 1) + rng.random(m)).astype(np.float32)
    got, window = lib.block_chr(v, max_block)
    exp, exp_window = orc.block_chr(v, max_block)
    assert got.tolist() == exp.tolist() and window == exp_window
    # a partition of [0, m): contiguous, in order, complete
    assert got[0, 0] == 0 and got[-1, 1] == m - 1
    assert (got[1:, 0] == got[:-1, 1] + 1).all() and (got[:, 1] >= got[:, 0]).all()


def test_host_block_chr_nan_row_sums(lib, orc):
    # a monomorphic marker gives NaN correlations (0/0 in tau_b), hence NaN row sums: comparisons with NaN are
    # false, so no cut can be placed next to it -- same in the oracle
    rng = np.random.default_rng(5)
    v = (4 + rng.random(400)).astype(np.float32)
    v[123] = np.nan
    got, _ = lib.block_chr(v, 60)
    exp, _ = orc.block_chr(v, 60)
    assert got.tolist() == exp.tolist()


def test_block_abi_argument_errors(lib):
    first = np.zeros(4, np.int64)
    assert lib.load_library().cusk_block_chr(None, 4, 10, first.ctypes.data_as(C.POINTER(C.c_int64)),
                                  first.ctypes.data_as(C.POINTER(C.c_int64)), None) < 0
    assert lib.load_library().cusk_hanning_smoothing(None, 4, 3, None) != 0
    assert lib.load_library().cusk_mcorrk_banded(None, None, 4, 4, 2, C.c_float(0), None, None, None) != 0
    assert lib.load_library().cusk_make_blocks(None, b"x", 10, C.c_float(0), 2) != 0


# ---- bit-plane routines of the banded kernel, compiled for the host ------------------------------------------

@pytest.fixture(scope="module")
def hostcore():
    out = os.path.join(ROOT, "tests", "_build", "libbanded_hostcheck.so")
    os.makedirs(os.path.dirname(out), exist_ok=True)
    subprocess.check_call(["g++", "-O2", "-std=c++17", "-shared", "-fPIC",
                           os.path.join(ROOT, "tests", "banded_host_core.cpp"), "-o", out])
    return C.CDLL(out)


@pytest.mark.parametrize("n", [1, 3, 4, 15, 16, 17, 31, 32, 33, 63, 64, 65, 100, 257, 1000])
def test_plane_tables_match_oracle(hostcore, orc, n):
    rng = np.random.default_rng(n)
    m = 6
    bed = random_bed(rng, m, n, miss_rate=0.1, miss_markers={1, 4})
    s9 = np.zeros(9, np.int64)
    used = C.c_int(0)
    seen = set()
    for a in range(m):
        for b in range(a + 1, m):
            exp = orc.pair_table(bed, n, a, b)
            for force in (0, 1):
                hostcore.bandc_host_pair_table(bed.ctypes.data_as(C.POINTER(C.c_ubyte)), C.c_size_t(n), C.c_size_t(a),
                                               C.c_size_t(b), force, s9.ctypes.data_as(C.POINTER(C.c_longlong)),
                                               C.byref(used))
                assert s9.tolist() == list(exp), (n, a, b, force)
                seen.add(used.value)
    if n >= 15:
        assert seen == {0, 1}  # both the four-count and the nine-count route were exercised


# ---- tile list of the tensor-core route (CUSK_BLOCK_TC=1): every band pair lies in exactly one listed tile ------

@pytest.mark.parametrize("Mc,R,W,MI", [
    (7, 7, 3, 64), (7, 7, 3, 42), (200, 200, 40, 64), (500, 470, 30, 42), (1000, 488, 512, 64), (1000, 488, 512, 42),
    (129, 129, 128, 64), (4096, 4096, 1, 64), (85, 43, 42, 42), (3000, 2000, 1000, 64),
])
def test_gram_band_tiles_cover_the_band_once(hostcore, Mc, R, W, MI):
    hostcore.bandc_host_gram_tiles.restype = C.c_size_t
    hostcore.bandc_host_gram_tiles.argtypes = [C.c_size_t] * 4 + [C.POINTER(C.c_int), C.c_size_t]
    cap = Mc * (W // 84 + 4)  # the capacity cusk_mcorrk_banded reserves
    buf = np.zeros(2 * cap, np.int32)
    cnt = hostcore.bandc_host_gram_tiles(Mc, R, W, MI, buf.ctypes.data_as(C.POINTER(C.c_int)), cap)
    assert 0 < cnt <= cap
    assert hostcore.bandc_host_gram_tiles(Mc, R, W, MI, None, 0) == cnt  # counting pass
    tiles = buf[:2 * cnt].reshape(cnt, 2)
    groups = (Mc + MI - 1) // MI
    assert (tiles[:, 0] >= 0).all() and (tiles[:, 0] < groups).all()
    assert (tiles[:, 1] >= 0).all() and (tiles[:, 1] < (groups + 1) // 2).all()
    listed = set(map(tuple, tiles.tolist()))
    assert len(listed) == cnt  # no tile twice
    i = np.arange(R)[:, None]
    j = i + 1 + np.arange(W)[None, :]
    ok = j < Mc
    need = set(zip((np.broadcast_to(i, j.shape)[ok] // MI).tolist(), (j[ok] // (2 * MI)).tolist()))
    assert need <= listed
    # and the list is tight: a listed tile is at most one block beyond what the band needs
    assert len(listed) <= len(need) + groups
    # the epilogue's filter (corr_gram.cu: j > i, j - i <= band_w, i < band_rows, both < m) applied to the listed
    # tiles writes every cell of the band exactly once and nothing else
    hits = np.zeros((R, W), np.int32)
    for tm, tn in tiles.tolist():
        ii = np.arange(tm * MI, min((tm + 1) * MI, Mc))[:, None]
        jj = np.arange(2 * tn * MI, min((2 * tn + 2) * MI, Mc))[None, :]
        keep = (jj > ii) & (jj - ii <= W) & (ii < R)
        a, b = np.nonzero(keep)
        np.add.at(hits, (ii[a, 0], jj[0, b] - ii[a, 0] - 1), 1)
    assert np.array_equal(hits, ok.astype(np.int32))


# ---- the `mps block` process without a device: argument handling, and no CPU route --------------------------

def test_mps_block_arguments_and_no_cpu_fallback(tmp_path):
    import cigwas_b200

    mps = cigwas_b200.mps_path()
    r = subprocess.run([mps, "block", "x", "10"], capture_output=True, text=True)
    assert r.returncode == 1 and "usage: mps block <bfiles> <max-block-size> <device-mem-gb> <corr-width>" in r.stdout
    r = subprocess.run([mps, "block", str(tmp_path / "absent"), "10", "1", "5"], capture_output=True, text=True)
    assert r.returncode == 1 and "file or directory not found" in r.stdout  # check_path, io.cpp:399-406
    # a valid file set: without a CUDA device the command must fail, not compute on the host
    rng = np.random.default_rng(1)
    n, m = 40, 30
    bed = random_bed(rng, m, n)
    stem = str(tmp_path / "g")
    with open(stem + ".bed", "wb") as f:
        f.write(bytes([0x6C, 0x1B, 0x01]) + bed.tobytes())
    with open(stem + ".bim", "w") as f:
        f.writelines(f"1\trs{i}\t0\t{i + 1}\tA\tG\n" for i in range(m))
    with open(stem + ".fam", "w") as f:
        f.writelines(f"f{i} i{i} 0 0 0 -9\n" for i in range(n))
    have_gpu = subprocess.run(["nvidia-smi", "-L"], capture_output=True).returncode == 0 if _which("nvidia-smi") else False
    r = subprocess.run([mps, "block", stem, "10", "1", "5"], capture_output=True, text=True)
    if have_gpu:
        assert r.returncode == 0 and os.path.exists(stem + "_m10.blocks")
    else:
        assert r.returncode != 0 and not os.path.exists(stem + "_m10.blocks")


def _which(name):
    import shutil

    return shutil.which(name)
