"""GPU parity tests of `sepselect`: sepselect_pairs_kernel through the C ABI and the `mps sepselect`
process against golden outputs of the reference's Python (cusk_postprocessing/sepselect.py, made by
tests/golden/make_sepselect_golden.py), and a 40 000-pair batch against the device routine compiled for the host
(tests/sepselect_host_core.cpp, test infrastructure).  Bars: see tests/sepselect_check.py."""
import ctypes as C
import os
import subprocess

import numpy as np
import pytest

from tests import sepselect_check as sc

pytestmark = pytest.mark.gpu
ROOT = sc.ROOT


@pytest.fixture(scope="module")
def cases():
    return sc.load_cases()


@pytest.fixture(scope="module")
def ctx():
    import cigwas_b200

    c = cigwas_b200.Context(0)
    yield c
    c.close()


@pytest.mark.parametrize("name", sc.CASES)
def test_sepselect_matches_reference(ctx, cases, tmp_path, name):
    case = cases[name]
    stem, prior = sc.unpack_inputs(case, tmp_path)
    ctx.sepselect(stem, case["info"]["alpha"], case["info"]["num_samples"], str(tmp_path / "max_sep_min_pc"), prior)
    sc.compare_outputs(case, tmp_path)


def test_mps_sepselect_process(cases, tmp_path):
    import cigwas_b200

    case = cases["prior"]
    stem, prior = sc.unpack_inputs(case, tmp_path)
    env = dict(os.environ, CUSK_DEVICES="0")
    r = subprocess.run([cigwas_b200.mps_path(), "sepselect", stem, repr(case["info"]["alpha"]), str(case["info"]["num_samples"]), prior],
                       capture_output=True, text=True, env=env)
    assert r.returncode == 0, r.stdout + r.stderr
    assert "Sepselect / v-structs done." in r.stdout  # ci-gwas.py:476
    sc.compare_outputs(case, tmp_path)  # default output stem: <dirname of the stem>/max_sep_min_pc (ci-gwas.py:475)


def test_pair_batch_matches_host_compiled_device_routine(ctx, tmp_path):
    """40 000 outer pairs over 17 traits and 3 000 variables: every field of every result equal to the same routine
    compiled for the host (sepselect.cu is built with -fmad=false, the host copy with -ffp-contract=off: the same
    operations in the same order; only log() comes from different libraries)"""
    so = str(tmp_path / "libhostcheck.so")
    subprocess.check_call(["g++", "-O2", "-std=c++17", "-ffp-contract=off", "-shared", "-fPIC",
                           os.path.join(ROOT, "tests", "sepselect_host_core.cpp"), "-o", so])
    host = C.CDLL(so)
    rng = np.random.default_rng(20261104)
    P, n, npairs = 17, 3000, 54000
    F = rng.standard_normal((n, 40))
    Cm = F @ F.T + 20.0 * np.eye(n)
    d = np.sqrt(np.diag(Cm))
    Cm = Cm / d[:, None] / d[None, :]
    Ct = np.ascontiguousarray(Cm[: P + 1])
    pairs = np.stack([rng.integers(0, n, npairs), rng.integers(0, P, npairs)], axis=1).astype(np.int32)
    swap = rng.random(npairs) < 0.5
    pairs[swap] = pairs[swap][:, ::-1]
    tmask = rng.integers(0, 1 << P, n, dtype=np.uint32)
    tmask[:P] &= ~(np.uint32(1) << np.arange(P, dtype=np.uint32))  # no variable is its own neighbour
    # outer pairs of unshielded triples: j is not adjacent to i (with an end of the pair among the candidates the
    # Schur complement is singular and rounding noise decides; the C ABI rejects such a pair, tested below)
    i, j = pairs[:, 0], pairs[:, 1]
    pairs = pairs[(i != j) & ~((j < P) & ((tmask[i] >> np.minimum(j, 31).astype(np.uint32)) & 1).astype(bool))]
    thr = 3.89 / np.sqrt(20000.0 - np.arange(P + 1) - 3)
    ln, fm, st = ctx.sepselect_pairs(Ct, P, pairs, tmask, thr)
    want = np.zeros((pairs.shape[0], 34), np.int32)
    host.cusk_sepselect_pairs.argtypes = [C.c_void_p, C.c_void_p, C.c_int64, C.c_int, C.c_int, C.c_void_p, C.c_int64, C.c_void_p,
                                          C.c_void_p, C.c_void_p]
    assert host.cusk_sepselect_pairs(None, Ct.ctypes.data, n, P, n, pairs.ctypes.data, pairs.shape[0], tmask.ctypes.data,
                                     thr.ctypes.data, want.ctypes.data) == 0
    differ = np.flatnonzero((want[:, 0] != ln) | (want[:, 1] != fm) | (want[:, 2:] != st).any(axis=1))
    # log() of the two maths libraries may differ in the last place: allow a decision to flip on a razor-edge comparison
    assert differ.size <= pairs.shape[0] // 2000, (differ.size, differ[:10])
    assert pairs.shape[0] > 39000 and ln.max() > 3 and (ln == 0).any()
    # an end of the pair among its own candidates: rejected (CUSK_E_ARG), not searched
    for bad_pair, var, bit in (((0, P + 5), 0, 0), ((P + 7, 3), P + 7, 3)):
        bad_mask = tmask.copy()
        bad_mask[var] |= np.uint32(1 << bit)
        with pytest.raises(Exception, match="among the trait neighbours"):
            ctx.sepselect_pairs(Ct, P, np.array([bad_pair], np.int32), bad_mask, thr)
