#!/usr/bin/env python3
"""Extract the reference's own golden vectors into tests/golden/golden.json.

Run ONCE in the build container (where /root/reference is mounted); the output is committed
so that nothing in tests/, smoke() or bench.py needs /root/reference at run time.

Sources (relative to /root/reference/cusk):
  include/test_data/cupc_test_set.h        C_N10 / A_N10 skeleton pair (tests/cupc_tests.cpp:17-88)
  include/test_data/bed_marker_test_set.h  10x3x2 and 100x7x5 BED correlation sets
  include/test_data/parent_set_test_set.h  20-node variable-selection graph (tests/parents_tests.cpp)
  tests/corr_tests.cpp                     inline expected values (:187-289) and unranking (:101-139)
  tests/prep_tests.cpp                     small.bed means/stds (:44-73)
  tests/cuskss_tests.cpp                   file-level cuskss expectations (:31-209)
  tests/test_files/*                       the small input files those tests read
"""
import base64
import json
import os
import re
import sys

REF = os.environ.get("CUSK_REFERENCE", "/root/reference/cusk")
OUT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden.json")


def strip_comments(src):
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return re.sub(r"//[^\n]*", "", src)


def c_array(src, name):
    """Return the numbers of `name[...] = {...}` / `name = {...}` / `name{...}` as floats."""
    m = re.search(r"\b" + re.escape(name) + r"\s*(\[[^\]]*\])?\s*=?\s*\{(.*?)\}\s*;", src, flags=re.S)
    if not m:
        raise KeyError(name)
    body = m.group(2)
    out = []
    for tok in body.replace("\n", " ").split(","):
        tok = tok.strip()
        if not tok:
            continue
        tok = tok.rstrip("fF") if not tok.lower().startswith("0x") else tok
        out.append(int(tok, 16) if tok.lower().startswith("0x") else float(tok))
    return out


def read(path):
    with open(os.path.join(REF, path), "r") as f:
        return strip_comments(f.read())


def b64(path):
    with open(os.path.join(REF, path), "rb") as f:
        return base64.b64encode(f.read()).decode()


def main():
    g = {}
    cupc = read("include/test_data/cupc_test_set.h")
    g["cupc_n10"] = {
        "alpha": 0.00001,
        "n": 10,
        "sample_size": 10000,
        "C": c_array(cupc, "C_N10"),
        "A": [int(v) for v in c_array(cupc, "A_N10")],
        "cite": "include/test_data/cupc_test_set.h:5-35; tests/cupc_tests.cpp:17-88",
    }
    g["threshold"] = {"n": 500000, "alpha": 0.00000001, "th0": 0.0081045, "tol": 0.0001,
                      "cite": "tests/cupc_tests.cpp:10-15"}

    bmt = read("include/test_data/bed_marker_test_set.h")
    g["bmt"] = {
        "n": 10, "m": 3, "p": 2,
        "bed": [int(v) for v in c_array(bmt, "bmt_marker_vals")],
        "mean": c_array(bmt, "bmt_marker_mean"),
        "std": c_array(bmt, "bmt_marker_std"),
        "phen": c_array(bmt, "bmt_phen_vals"),
        # tests/corr_tests.cpp:211-289
        "mxp_expected": [-0.982847, -0.856312, -0.380668, -0.405905, -0.695899, -0.928034],
        "mxm_npn_expected": [0.299969, 0.924167, 0.0],
        "pxp_expected": [0.808083],
        "tol": 0.00001,
    }
    g["bmt2"] = {
        "n": 100, "m": 7, "p": 5,
        "bed": [int(v) for v in c_array(bmt, "bmt2_marker_vals")],
        "mean": c_array(bmt, "bmt2_marker_mean"),
        "std": c_array(bmt, "bmt2_marker_std"),
        "phen": c_array(bmt, "bmt2_phen_vals"),
        "mxm_npn_expected": c_array(bmt, "bmt2_marker_corrs"),
        "mxm_pearson_expected": c_array(bmt, "bmt2_marker_corrs_pearson"),
        "mxp_expected": c_array(bmt, "bmt2_marker_phen_corrs_pearson"),
        "pxp_expected": c_array(bmt, "bmt2_phen_corrs"),
        "tol": 0.00001,
        "cite": "include/test_data/bed_marker_test_set.h:32-186; tests/corr_tests.cpp:141-185,291-336",
    }

    ps = read("include/test_data/parent_set_test_set.h")
    g["parent_set"] = {
        "num_markers": 17, "num_phen": 3,
        "adj": [int(v) for v in c_array(ps, "TEST_ADJ_MAT")],
        "d0": [int(v) for v in c_array(ps, "TEST_PAR_SET_D0")],
        "d1": [int(v) for v in c_array(ps, "TEST_PAR_SET_D1")],
        "d2": [int(v) for v in c_array(ps, "TEST_PAR_SET_D2")],
        "cite": "include/test_data/parent_set_test_set.h; tests/parents_tests.cpp:5-39",
    }

    # tests/prep_tests.cpp:44-73
    g["prep_small"] = {
        "stds": [0.66332496, 0.83066239, 0.6, 0.77459667, 0.83066239],
        "means": [0.6, 1.1, 0.8, 1.0, 0.9],
        "tol": 1e-6, "n": 10, "m": 5,
    }
    # tests/corr_tests.cpp:187-209 (pxp with NaN handling); expected values parsed from the test
    ct = read("tests/corr_tests.cpp")
    m = re.search(r"TEST\(cu_phen_corr_pearson_npn[^)]*\)(.*?)\nTEST\(", ct, flags=re.S)
    if m:
        try:
            g["with_nan_pxp_expected"] = c_array(m.group(1), "phen_corr_expected")
        except KeyError:
            pass
    # tests/corr_tests.cpp:101-139: upper-triangular unranking known answers
    g["unrank"] = [[0, 3, 0, 1], [1, 3, 0, 2], [2, 3, 1, 2], [28899002, 7603, 7601, 7602],
                   [28898997, 7603, 7599, 7600], [28898996, 7603, 7598, 7602],
                   [28898995, 7603, 7598, 7601]]  # [lin_ix, num_rows, row, col]

    # tests/cuskss_tests.cpp:31-209
    g["cuskss"] = {
        "alpha": 0.0001, "sample_size": 500000.0,
        "trait_only": {"adj": [0, 1, 1, 1, 0, 1, 1, 1, 0],
                       "corr": [1.0, 0.0608594558771734, 0.074239793758568, 0.0608594558771734, 1.0,
                                0.0675875270156859, 0.074239793758568, 0.0675875270156859, 1.0]},
        "two_stage": {"max_level": 3, "max_level_two": 1, "depth": 1, "ixs": [2, 3, 4, 5],
                      "adj": [0, 0, 0, 1, 0, 0, 1, 1, 0, 1, 0, 1, 1, 1, 1, 0],
                      "corr": [1.0, 0.0005, 0.0001, -0.01, 0.0005, 1.0, 0.0608594558771734,
                               0.074239793758568, 0.0001, 0.0608594558771734, 1.0, 0.0675875270156859,
                               -0.01, 0.074239793758568, 0.0675875270156859, 1.0]},
        "corr_tol": 0.001,
    }
    files = {}
    for fn in sorted(os.listdir(os.path.join(REF, "tests/test_files"))):
        files[fn] = b64("tests/test_files/" + fn)
    g["files_b64"] = files

    with open(OUT, "w") as f:
        json.dump(g, f, indent=1)
    print("wrote", OUT, os.path.getsize(OUT), "bytes")


if __name__ == "__main__":
    sys.exit(main())
