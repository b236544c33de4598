"""CPU checks of `sepselect` (cusk_postprocessing/sepselect.py:429-584) against golden outputs of the
reference's Python.  The product runs the greedy search on the GPU and has no CPU path, so here the device routine
(ci-gwas_b200/csrc/sepselect_core.h) is compiled for the host as TEST infrastructure (tests/sepselect_host_core.cpp)
and preloaded over the library's GPU entry point in a child process: that checks the host side (MatrixMarket input,
collinear markers, unshielded triples, ambiguous triples, orientation, output files) and the arithmetic of the device
routine.  The kernel launch itself is covered by tests/test_sepselect_gpu.py."""
import os
import subprocess
import sys

import pytest

from tests import sepselect_check as sc

ROOT = sc.ROOT
CHILD = r"""
import ctypes as C, os, sys
sys.path.insert(0, %(root)r)
import cigwas_b200
lib = cigwas_b200.load_library()
lib.cusk_sepselect.argtypes = [C.c_void_p, C.c_char_p, C.c_double, C.c_int64, C.c_char_p, C.c_char_p]
stem, alpha, ns, prior, out = sys.argv[1], float(sys.argv[2]), int(sys.argv[3]), sys.argv[4], sys.argv[5]
rc = lib.cusk_sepselect(C.c_void_p(1), stem.encode(), alpha, ns, None if prior == "NULL" else prior.encode(), out.encode())
if rc != 0:
    print(lib.cusk_last_error().decode())
sys.exit(0 if rc == 0 else 1)
"""


@pytest.fixture(scope="module")
def hostcheck():
    import cigwas_b200

    if not os.path.exists(cigwas_b200.library_path()):
        cigwas_b200.build_library()
    out = os.path.join(ROOT, "tests", "_build", "libsepselect_hostcheck.so")
    os.makedirs(os.path.dirname(out), exist_ok=True)
    subprocess.check_call(["g++", "-O2", "-std=c++17", "-shared", "-fPIC", os.path.join(ROOT, "tests", "sepselect_host_core.cpp"),
                           "-o", out])
    return out


@pytest.fixture(scope="module")
def cases():
    return sc.load_cases()


def run_child(hostcheck, stem, alpha, ns, prior, out):
    env = dict(os.environ, LD_PRELOAD=hostcheck)
    return subprocess.run([sys.executable, "-c", CHILD % {"root": ROOT}, stem, repr(alpha), str(ns), prior or "NULL", out],
                          capture_output=True, text=True, env=env)


@pytest.mark.parametrize("name", sc.CASES)
def test_host_side_and_device_routine_match_reference(hostcheck, cases, tmp_path, name):
    case = cases[name]
    stem, prior = sc.unpack_inputs(case, tmp_path)
    r = run_child(hostcheck, stem, case["info"]["alpha"], case["info"]["num_samples"], prior, str(tmp_path / "max_sep_min_pc"))
    assert r.returncode == 0, r.stdout + r.stderr
    if name == "collinear":
        assert "Removed 4 collinear markers" in r.stdout  # sepselect.py:483
    sc.compare_outputs(case, tmp_path)


def test_no_cpu_path_in_the_product(cases, tmp_path):
    import ctypes as C

    import cigwas_b200

    lib = cigwas_b200.load_library()
    lib.cusk_sepselect.argtypes = [C.c_void_p, C.c_char_p, C.c_double, C.c_int64, C.c_char_p, C.c_char_p]
    stem, _ = sc.unpack_inputs(cases["small"], tmp_path)
    rc = lib.cusk_sepselect(None, stem.encode(), 1e-4, 20000, None, str(tmp_path / "o").encode())
    assert rc == -4 and b"no CPU path" in lib.cusk_last_error()  # CUSK_E_NODEVICE
    assert not os.path.exists(tmp_path / "o.mdim")


def test_input_errors(hostcheck, cases, tmp_path):
    stem, _ = sc.unpack_inputs(cases["small"], tmp_path)
    r = run_child(hostcheck, str(tmp_path / "absent"), 1e-4, 20000, None, str(tmp_path / "o"))
    assert r.returncode == 1 and "absent.mdim" in r.stdout
    bad = tmp_path / "prior.bin"
    bad.write_bytes(b"\0" * 12)
    r = run_child(hostcheck, stem, 1e-4, 20000, str(bad), str(tmp_path / "o"))
    assert r.returncode == 1 and "n_trait * n_trait" in r.stdout  # sepselect.py:465
