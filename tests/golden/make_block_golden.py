#!/usr/bin/env python3
"""Extract the reference's golden vectors for `mps block` into tests/golden/block_golden.json.

Run ONCE in the build container (where /root/reference is mounted); the output is committed.

Sources (relative to /root/reference/cusk):
  include/test_data/blocking_test_set.h    TEST_V (3 998 row sums), TEST_V_SMOOTH (hanning_smoothing of the first
                                           1 000 with window 101; tests/blocking_tests.cpp:40-54, tolerance 0.01)
  tests/blocking_tests.cpp:9-38            block_chr(TEST_V, "1", 500): 17 expected blocks (exact)
  include/test_data/bed_marker_test_set.h  bmt2_sparse_mcorrs: cal_mcorrk_banded of the 100 x 7 BED set with
                                           corr_width 3 (tests/corr_tests.cpp:62-75, tolerance 1e-5); the BED bytes
                                           themselves are already in golden.json ("bmt2")
"""
import json
import os
import re

from make_golden import REF, c_array, read

OUT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "block_golden.json")


def main():
    g = {}
    bts = read("include/test_data/blocking_test_set.h")
    g["test_v"] = c_array(bts, "TEST_V")
    g["test_v_smooth"] = c_array(bts, "TEST_V_SMOOTH")
    g["smooth_n"], g["smooth_window"], g["smooth_tol"] = 1000, 101, 0.01
    tests = read("tests/blocking_tests.cpp")
    body = tests[tests.index("TEST(block_chr, expected_results_synthetic_data)"):tests.index("TEST(hanning_smoothing")]
    g["block_chr"] = {
        "chr": "1",
        "max_block_size": int(re.search(r'block_chr\(TEST_V,\s*"1",\s*(\d+)\)', body).group(1)),
        "blocks": [[int(a), int(b)] for a, b in re.findall(r'MarkerBlock\("1",\s*(\d+),\s*(\d+),\s*0\)', body)],
        "cite": "tests/blocking_tests.cpp:9-38",
    }
    bm = read("include/test_data/bed_marker_test_set.h")
    g["bmt2_sparse_mcorrs"] = {"corr_width": 3, "expected": c_array(bm, "bmt2_sparse_mcorrs"), "tol": 1e-5,
                               "cite": "tests/corr_tests.cpp:62-75"}
    assert len(g["test_v"]) == 3998 and len(g["test_v_smooth"]) == 1000 and len(g["block_chr"]["blocks"]) == 17
    assert len(g["bmt2_sparse_mcorrs"]["expected"]) == 21
    with open(OUT, "w") as f:
        json.dump(g, f)
    print("wrote", OUT, os.path.getsize(OUT), "bytes")


if __name__ == "__main__":
    main()
