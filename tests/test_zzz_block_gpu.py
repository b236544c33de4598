"""GPU parity tests of `mps block`: the banded correlation kernels through the C ABI
(cusk_mcorrk_banded) against the reference's golden vector (cal_mcorrk_banded on the 100 x 7 set, tolerance 1e-5,
cusk/tests/corr_tests.cpp:62-75) and against the oracle on seeded genotypes (with and without missing calls, sample
counts that are not multiples of 4 / 32 / 1024, chunked runs under a device-memory bound), and the `mps block`
process against the oracle's blocks.

Bars: correlations within 2e-6 of the oracle (both evaluate the same FP64 formula on the same exact integer table;
device and host sqrt/sin may differ in the last place before the cast to float); row sums equal to the float
left-to-right sum of the device's own correlations (bit-exact); blocks identical to the oracle's."""
import json
import os
import subprocess

import numpy as np
import pytest

from tests.test_block_cpu import random_bed

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
TOL = 2e-6


@pytest.fixture(scope="module")
def ctx():
    import cigwas_b200

    c = cigwas_b200.Context(0)
    yield c
    c.close()


def seq_abs_sums(corrs):
    s = np.zeros(corrs.shape[0], np.float32)
    for k in range(corrs.shape[1]):
        s = (s + np.abs(corrs[:, k])).astype(np.float32)
    return s


def check_against_oracle(ctx, oracle, bed, m, n, w, **kw):
    r = ctx.mcorrk_banded(bed, m, n, w, **kw)
    exp = oracle.mcorrk_banded(bed, m, n, w)
    got = r["corrs"]
    assert np.array_equal(np.isnan(got), np.isnan(exp))
    d = np.nanmax(np.abs(got - exp)) if np.isfinite(exp).any() else 0.0
    assert d <= TOL, (m, n, w, float(d), np.argwhere(np.abs(got - exp) > TOL)[:5].tolist())
    # structural zeros beyond the last marker
    for row in range(max(0, m - w), m):
        assert (got[row, m - 1 - row:] == 0).all()
    sums = r["row_abs_sums"]
    ref = seq_abs_sums(got)
    assert np.array_equal(sums, ref, equal_nan=True), float(np.nanmax(np.abs(sums - ref)))
    return r


def test_banded_golden(ctx, golden):
    with open(os.path.join(ROOT, "tests", "golden", "block_golden.json")) as f:
        g = json.load(f)["bmt2_sparse_mcorrs"]
    b = golden["bmt2"]
    r = ctx.mcorrk_banded(np.array(b["bed"], np.uint8), b["m"], b["n"], g["corr_width"])
    d = np.abs(r["corrs"].ravel() - np.array(g["expected"])).max()
    assert d < g["tol"], float(d)
    assert r["stats"]["pairs"] == 4 * 3 + 3  # rows 0..3 full, then 2 + 1 + 0


def test_reference_named_entry_point(golden):
    """cu_marker_corr_pearson_npn_batched_sparse with the reference's own arguments (corr_tests.cpp:76-99: corr_width 3,
    batch_size 5) -- the batch size only bounded the reference's device buffers"""
    import ctypes as C

    import cigwas_b200

    lib = cigwas_b200.load_library()
    with open(os.path.join(ROOT, "tests", "golden", "block_golden.json")) as f:
        g = json.load(f)["bmt2_sparse_mcorrs"]
    b = golden["bmt2"]
    bed = np.array(b["bed"], np.uint8)
    out = np.zeros(3 * b["m"], np.float32)
    lib.cu_marker_corr_pearson_npn_batched_sparse.restype = None
    lib.cu_marker_corr_pearson_npn_batched_sparse.argtypes = [C.POINTER(C.c_ubyte), C.c_size_t, C.c_size_t, C.c_size_t,
                                                              C.c_size_t, C.POINTER(C.c_float)]
    lib.cu_marker_corr_pearson_npn_batched_sparse(bed.ctypes.data_as(C.POINTER(C.c_ubyte)), b["m"], b["n"], 3, 5,
                                                  out.ctypes.data_as(C.POINTER(C.c_float)))
    assert np.abs(out - np.array(g["expected"])).max() < g["tol"]


@pytest.mark.parametrize("m,n,w,miss", [
    (7, 100, 3, 0.0), (40, 33, 5, 0.0), (40, 33, 39, 0.05), (70, 1023, 31, 0.0), (70, 1025, 32, 0.0),
    (100, 2050, 33, 0.02), (130, 4097, 64, 0.0), (97, 777, 96, 0.0), (300, 1500, 20, 0.0),
])
def test_banded_matches_oracle(ctx, oracle, m, n, w, miss):
    rng = np.random.default_rng(m * 1000 + n)
    bed = random_bed(rng, m, n, miss_rate=miss)
    check_against_oracle(ctx, oracle, bed, m, n, w)


def test_banded_mixed_tiles(ctx, oracle):
    """missing calls in a few markers only: tiles that touch them take the nine-count route, the rest the four-count
    route, inside one launch"""
    rng = np.random.default_rng(77)
    m, n, w = 200, 3000, 40
    bed = random_bed(rng, m, n, miss_rate=0.03, miss_markers={5, 70, 71, 199})
    check_against_oracle(ctx, oracle, bed, m, n, w)


def test_banded_monomorphic_marker_gives_nan(ctx, oracle):
    rng = np.random.default_rng(3)
    m, n, w = 40, 500, 6
    bed = random_bed(rng, m, n)
    bed[11] = 0xFF  # every sample genotype 0
    r = check_against_oracle(ctx, oracle, bed, m, n, w)
    assert np.isnan(r["corrs"][11, :w]).all() and np.isnan(r["row_abs_sums"][11])


def test_banded_chunked_equals_whole(ctx):
    """a device-memory bound that forces several chunks of rows (each re-loading the corr_width markers to its
    right) changes nothing"""
    rng = np.random.default_rng(11)
    m, n, w = 500, 4000, 30
    bed = random_bed(rng, m, n, miss_rate=0.01, miss_markers={100, 333})
    whole = ctx.mcorrk_banded(bed, m, n, w)
    per_marker = (n + 3) // 4 + 12 * 128 + 16 + 4 * w + 4
    part = ctx.mcorrk_banded(bed, m, n, w, device_mem_gb=per_marker * 120 / 1e9)
    assert part["stats"]["chunks"] >= 5 and whole["stats"]["chunks"] == 1
    assert np.array_equal(whole["corrs"], part["corrs"], equal_nan=True)
    assert np.array_equal(whole["row_abs_sums"], part["row_abs_sums"], equal_nan=True)
    import cigwas_b200

    with pytest.raises(cigwas_b200.CuskError):  # batch <= corr width: the reference exits 1 here (corr_host.cu:82-91)
        ctx.mcorrk_banded(bed, m, n, w, device_mem_gb=per_marker * 20 / 1e9)
    with pytest.raises(cigwas_b200.CuskError):
        ctx.mcorrk_banded(bed, m, n, m)


@pytest.mark.parametrize("m,n,w,miss,mem_markers", [
    (7, 100, 3, 0.0, 0), (130, 4097, 64, 0.0, 0), (300, 1500, 20, 0.02, 0), (97, 777, 96, 0.0, 0),
    (700, 5000, 150, 0.0, 0), (700, 5000, 150, 0.01, 0), (900, 3000, 40, 0.005, 200),
])
def test_banded_tensor_core_route_equals_default(ctx, oracle, m, n, w, miss, mem_markers):
    """CUSK_BLOCK_TC=1 (tcgen05 indicator Gram over the band's tiles, the default from 512 markers on) against
    CUSK_BLOCK_TC=0 (AND + POPC bit planes): both
    build the same exact integer tables and evaluate the same FP64 expression -> the same floats"""
    rng = np.random.default_rng(m + n + w)
    bed = random_bed(rng, m, n, miss_rate=miss, miss_markers={3, m // 2} if miss else None)
    kw = {}
    if mem_markers:  # a device-memory bound that forces chunks on the tensor-core route too
        kp = (n + 127) // 128 * 128
        per_marker = (n + 3) // 4 + (128 * kp + 41) // 42 + 16 + 8 * (w // 84 + 4) + 4 * w + 4
        kw["device_mem_gb"] = (per_marker * mem_markers + 384 * kp) / 1e9
    os.environ["CUSK_BLOCK_TC"] = "0"
    try:
        base = ctx.mcorrk_banded(bed, m, n, w)
        os.environ["CUSK_BLOCK_TC"] = "1"
        tc = check_against_oracle(ctx, oracle, bed, m, n, w, **kw)
    finally:
        del os.environ["CUSK_BLOCK_TC"]
    if mem_markers:
        assert tc["stats"]["chunks"] > 1
    assert np.array_equal(base["corrs"], tc["corrs"], equal_nan=True)
    assert np.array_equal(base["row_abs_sums"], tc["row_abs_sums"], equal_nan=True)


def write_bfiles(tmp_path, beds, n):
    """beds: list of (chr id, packed columns [m, ceil(n/4)])"""
    stem = str(tmp_path / "geno")
    with open(stem + ".bed", "wb") as f:
        f.write(bytes([0x6C, 0x1B, 0x01]))
        for _, b in beds:
            f.write(b.tobytes())
    with open(stem + ".bim", "w") as f:
        for cid, b in beds:
            for i in range(b.shape[0]):
                f.write(f"{cid}\trs{cid}_{i}\t0\t{i + 1}\tA\tG\n")
    with open(stem + ".fam", "w") as f:
        for i in range(n):
            f.write(f"f{i} i{i} 0 0 0 -9\n")
    return stem


def test_mps_block_process(oracle, tmp_path):
    import cigwas_b200

    rng = np.random.default_rng(20261104)
    n, w, max_block = 600, 25, 150
    beds = [("1", random_bed(rng, 700, n, ld=0.97)), ("2", random_bed(rng, 400, n, ld=0.97, miss_rate=0.01))]
    stem = write_bfiles(tmp_path, beds, n)
    env = dict(os.environ, CUSK_DEVICES="0")
    r = subprocess.run([cigwas_b200.mps_path(), "block", stem, str(max_block), "1", str(w)], capture_output=True, text=True,
                       env=env)
    assert r.returncode == 0, r.stdout + r.stderr
    with open(f"{stem}_m{max_block}.blocks") as f:
        lines = [ln.rstrip("\n").split("\t") for ln in f]
    exp = []
    for cid, b in beds:
        sums = oracle.banded_row_abs_sums(oracle.mcorrk_banded(b, b.shape[0], n, w))
        blocks, _ = oracle.block_chr(sums, max_block)
        exp += [[cid, str(a), str(z)] for a, z in blocks.tolist()]
    assert lines == exp
    # a second run APPENDS (io.cpp:269 opens with ios::app)
    r = subprocess.run([cigwas_b200.mps_path(), "block", stem, str(max_block), "1", str(w)], capture_output=True, text=True,
                       env=env)
    assert r.returncode == 0
    with open(f"{stem}_m{max_block}.blocks") as f:
        assert len(f.readlines()) == 2 * len(exp)
    # corr width >= markers of a chromosome: exit status 1
    r = subprocess.run([cigwas_b200.mps_path(), "block", stem, str(max_block), "1", "400"], capture_output=True, text=True,
                       env=env)
    assert r.returncode == 1
