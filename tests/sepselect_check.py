"""Shared by the CPU and GPU tests of `sepselect`: unpack a golden case (tests/golden/sepselect_golden.json, made by
running cusk_postprocessing/sepselect.py, see tests/golden/make_sepselect_golden.py) and compare a run's outputs.

The bar: .mdim and .ssm (maximal separating sets, in the order the traits were added) byte-identical; .ut / .atr equal
as sets of triples (the reference writes them in Python set order); the three MatrixMarket files equal as matrices
(exact, also the correlations: both writers print round-trip digits)."""
import base64
import io
import json
import os

import numpy as np
from scipy.io import mmread

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CASES = ["small", "traits17", "collinear", "prior"]


def load_cases():
    with open(os.path.join(ROOT, "tests", "golden", "sepselect_golden.json")) as f:
        return json.load(f)


def unpack_inputs(case, d):
    for fn, b64 in case["in"].items():
        with open(os.path.join(d, fn), "wb") as f:
            f.write(base64.b64decode(b64))
    prior = os.path.join(d, "prior.bin")
    return os.path.join(d, "merged_blocks"), (prior if case["info"]["prior"] else None)


def triples(raw):
    a = np.frombuffer(raw, np.int32).reshape(-1, 3)
    return sorted(map(tuple, a.tolist()))


def dense(raw):
    return mmread(io.BytesIO(raw), spmatrix=False).toarray()


def compare_outputs(case, d, stem="max_sep_min_pc"):
    want = {k: base64.b64decode(v) for k, v in case["out"].items()}
    got = {}
    for fn in want:
        with open(os.path.join(d, fn.replace("max_sep_min_pc", stem)), "rb") as f:
            got[fn] = f.read()
    assert got["max_sep_min_pc.mdim"] == want["max_sep_min_pc.mdim"]
    assert got["max_sep_min_pc.ssm"] == want["max_sep_min_pc.ssm"]
    for sfx in (".ut", ".atr"):
        assert triples(got["max_sep_min_pc" + sfx]) == triples(want["max_sep_min_pc" + sfx]), sfx
    for sfx in ("_sam.mtx", "_spm.mtx", "_scm.mtx"):
        g, w = dense(got["max_sep_min_pc" + sfx]), dense(want["max_sep_min_pc" + sfx])
        assert g.shape == w.shape and np.array_equal(g, w, equal_nan=True), sfx
    nset = sum(1 for v in case["info"]["max_sepset_len"].values() if v > 0)
    assert len(got["max_sep_min_pc.ssm"].splitlines()) == nset
