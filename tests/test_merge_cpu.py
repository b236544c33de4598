"""CPU parity tests of the native `merge-block-outputs` and cuskss-merged reformatting (ci-gwas_b200/host/merge.cpp)
against the reference's own Python (cusk_postprocessing/merge_blocks.py:300-425).  The expected files in
tests/golden/merge_golden.json were produced by running that Python on the same inputs
(tests/golden/make_merge_golden.py); the bar is byte identity of all four output files."""
import base64
import json
import os
import subprocess
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def cases():
    with open(os.path.join(ROOT, "tests", "golden", "merge_golden.json")) as f:
        return json.load(f)


@pytest.fixture(scope="module")
def pkg():
    import cigwas_b200

    if not os.path.exists(cigwas_b200.library_path()):
        cigwas_b200.build_library()
    cigwas_b200.load_library()
    return cigwas_b200


def unpack(files, d):
    for fn, b64 in files.items():
        with open(os.path.join(d, fn), "wb") as f:
            f.write(base64.b64decode(b64))


def check_outputs(expected, d):
    for fn, b64 in expected.items():
        with open(os.path.join(d, fn), "rb") as f:
            got = f.read()
        want = base64.b64decode(b64)
        assert got == want, f"{fn} differs from the reference's output:\n{got[:400]!r}\n{want[:400]!r}"


@pytest.mark.parametrize("name", ["four_blocks", "first_missing", "dense_random"])
def test_merge_block_outputs_matches_reference_bytes(cases, pkg, tmp_path, name):
    case = cases[name]
    unpack(case["in"], tmp_path)
    assert sorted(case["out"]) == ["merged_blocks.ixs", "merged_blocks.mdim", "merged_blocks_sam.mtx", "merged_blocks_scm.mtx"]
    pkg.merge_block_outputs(str(tmp_path / "all.blocks"), str(tmp_path))
    check_outputs(case["out"], tmp_path)


def test_mps_merge_process_contract(cases, pkg, tmp_path):
    case = cases["four_blocks"]
    unpack(case["in"], tmp_path)
    r = subprocess.run([pkg.mps_path(), "merge", str(tmp_path / "all.blocks"), str(tmp_path)], capture_output=True, text=True)
    assert r.returncode == 0, r.stdout + r.stderr
    assert f"Missing: {tmp_path}/1_10_29" in r.stdout  # merge_blocks.py:383
    check_outputs(case["out"], tmp_path)
    # through the command-line front end (ci-gwas.py:254-270: <cusk-output-dir> <blockfile>)
    for fn in case["out"]:
        os.remove(tmp_path / fn)
    r = subprocess.run([sys.executable, os.path.join(ROOT, "ci_gwas_b200.py"), "merge-block-outputs", str(tmp_path),
                        str(tmp_path / "all.blocks")], capture_output=True, text=True)
    assert r.returncode == 0, r.stdout + r.stderr
    check_outputs(case["out"], tmp_path)


def test_reformat_cuskss_merged_matches_reference_bytes(cases, pkg, tmp_path):
    case = cases["reformat"]
    unpack(case["in"], tmp_path)
    mdim = (tmp_path / "cuskss_merged.mdim").read_bytes()
    pkg.reformat_cuskss_merged(str(tmp_path))
    check_outputs(case["out"], tmp_path)
    assert (tmp_path / "cuskss_merged.mdim").read_bytes() == mdim
    # .ixs now holds the global (.bim row) indices of the retained markers, ascending
    assert np.fromfile(tmp_path / "cuskss_merged.ixs", np.int32).tolist() == [9, 12, 250, 251]


def test_errors(pkg, tmp_path):
    with pytest.raises(pkg.CuskError):
        pkg.merge_block_outputs(str(tmp_path / "absent.blocks"), str(tmp_path))
    (tmp_path / "b.blocks").write_text("1\t0\t9\n1\t10\t19\n")
    with pytest.raises(pkg.CuskError):  # no block has outputs (the reference dies with a NameError)
        pkg.merge_block_outputs(str(tmp_path / "b.blocks"), str(tmp_path))
    with pytest.raises(pkg.CuskError):
        pkg.reformat_cuskss_merged(str(tmp_path))
    r = subprocess.run([pkg.mps_path(), "merge", str(tmp_path / "b.blocks")], capture_output=True, text=True)
    assert r.returncode == 1 and "usage: mps merge" in r.stdout


def test_correlation_text_is_python_float_repr(pkg, tmp_path):
    """write_mm prints f"{numpy.float32}" = repr(float(x)): shortest round-trip digits of the double, exponent form
    below 1e-4 and from 1e16 (merge_blocks.py:322); checked here on values of every magnitude"""
    rng = np.random.default_rng(3)
    n = 40
    mant = rng.uniform(-1, 1, (n, n))
    expo = rng.integers(-44, 39, (n, n))
    with np.errstate(over="ignore"):
        c = (mant * 10.0 ** expo).astype(np.float32)
    c[0, :12] = [1.0, -1.0, 1e-4, 9.9999e-5, 1e16, 9.9999e15, 16777216.0, 0.1, 1e-45, 3.4028235e38, np.inf, -np.inf]
    c[1, :3] = [np.nan, -0.0, 0.5]
    (tmp_path / "9_0_99.mdim").write_text(f"{n}\t2\t14\n")
    c.tofile(tmp_path / "9_0_99.corr")
    a = np.zeros((n, n), np.int32)
    a[0, 1] = a[1, 0] = 1
    a.tofile(tmp_path / "9_0_99.adj")
    np.arange(n, dtype=np.int32).tofile(tmp_path / "9_0_99.ixs")
    (tmp_path / "b.blocks").write_text("9\t0\t99\n")
    pkg.merge_block_outputs(str(tmp_path / "b.blocks"), str(tmp_path))
    lines = (tmp_path / "merged_blocks_scm.mtx").read_text().splitlines()
    m = n - 2
    new = np.where(np.arange(n) < m, np.arange(n) + 2 + 1, np.arange(n) - m + 1)
    want = [f"{new[i]}\t{new[j]}\t{float(c[i, j])!r}" for i in range(n) for j in range(n) if c[i, j] != 0]
    assert lines[0] == "%%MatrixMarket matrix coordinate real general"
    assert lines[1] == f"{new[1]}\t{new[1]}\t{len(want)}"  # N = M = largest row index of the ADJACENCY entries
    assert lines[2:] == want
