"""CPU-only checks of the boundary: the C-ABI library builds for sm_100a, loads, exports every symbol
include/cusk_b200.h declares, and fails loudly (no fallback) when there is no CUDA device."""
import os
import re

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def lib():
    import cigwas_b200

    if not os.path.exists(cigwas_b200.library_path()):
        cigwas_b200.build_library()
    return cigwas_b200.load_library()


def declared_symbols():
    names = set()
    for fn in os.listdir(os.path.join(ROOT, "include")):
        if not fn.endswith(".h"):
            continue
        txt = open(os.path.join(ROOT, "include", fn)).read()
        txt = re.sub(r"/\*.*?\*/", "", txt, flags=re.S)
        for m in re.finditer(r"^[A-Za-z_][\w \*]*?[\s\*]([A-Za-z_]\w*)\s*\(", txt, flags=re.M):
            if m.group(1) not in ("defined",):
                names.add(m.group(1))
    return sorted(names)


def test_every_declared_symbol_is_exported(lib):
    syms = declared_symbols()
    assert "Skeleton" in syms and "hetcor_skeleton" in syms and "cusk_ctx_create" in syms
    missing = [s for s in syms if not hasattr(lib, s)]
    assert not missing, missing


def test_thresholds_cpu(lib, oracle, golden):
    import cigwas_b200

    t = golden["threshold"]
    assert abs(cigwas_b200.threshold_array(t["n"], t["alpha"])[0] - t["th0"]) < t["tol"]
    assert (cigwas_b200.threshold_array(10000, 1e-4) == oracle.threshold_array(10000, 1e-4)).all()


def test_no_cpu_fallback(lib):
    import torch

    import cigwas_b200

    if torch.cuda.is_available():
        pytest.skip("a CUDA device is present")
    with pytest.raises(cigwas_b200.CuskError):
        cigwas_b200.Context(0)


def test_product_does_not_import_oracle():
    """The product path must never route through oracle/ (tier rule 3)."""
    pkg = os.path.join(ROOT, "ci-gwas_b200")
    for dirpath, _, files in os.walk(pkg):
        for fn in files:
            if fn.endswith((".py", ".cu", ".cuh", ".cpp", ".h", ".hpp")):
                txt = open(os.path.join(dirpath, fn), errors="ignore").read()
                assert "oracle_lib" not in txt and "libcusk_oracle" not in txt and "orc_" not in txt, fn


def test_synth_is_deterministic():
    from cigwas_b200 import synth

    a = synth.sem_correlation(60, 3, stream=5, num_samples=1000, causal_per_trait=5)
    b = synth.sem_correlation(60, 3, stream=5, num_samples=1000, causal_per_trait=5)
    assert (a == b).all() and (a == a.T).all() and (np.diag(a) == 1).all()
