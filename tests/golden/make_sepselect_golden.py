#!/usr/bin/env python3
"""Golden input/output pairs for the separating-set selection + v-structure orientation step (`ci-gwas.py sepselect` /
`orient-v-structs`), made by RUNNING the reference's own Python (cusk_postprocessing/sepselect.py:
MergedCuskResults :429-467, rm_collinear_markers :469-483, find_maximal_and_min_pcorr_sepsets_incr :262-329,
mark_ambiguous_triples :95-107, orient_v_structures :485-511, to_file :545-573) on small merged skeletons.

Run ONCE in the build container (where /root/reference is mounted); the result, tests/golden/sepselect_golden.json, is
committed so that no test needs /root/reference at run time.  Files are stored base64-encoded.

The inputs are written by the reference's GlobalBdpcResult.write_mm (merge_blocks.py:300-325), i.e. they are exactly what
`merge-block-outputs` produces: traits first, 1-based MatrixMarket coordinates.

Cases:
  small          6 traits, 30 markers
  traits17       17 traits with a dense trait graph (deep greedy searches), 60 markers
  collinear      8 traits, 40 markers of which three pairs are exactly collinear (rm_collinear_markers), one marker
                 collinear with a trait
  prior          `small` with an orientation prior file over the traits
"""
import base64
import json
import os
import sys
import tempfile

import numpy as np

sys.path.insert(0, os.environ.get("CIGWAS_REFERENCE", "/root/reference"))
from cusk_postprocessing.merge_blocks import GlobalBdpcResult  # noqa: E402
from cusk_postprocessing.sepselect import orient_v_structures_merged  # noqa: E402

OUT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "sepselect_golden.json")


def make_inputs(d, P, M, seed, n_samples, trait_density=0.5, collinear=()):
    rng = np.random.default_rng(seed)
    n = P + M
    X = rng.standard_normal((n_samples, M))
    for k in range(1, M):
        if rng.random() < 0.3:
            X[:, k] = 0.6 * X[:, k - 1] + 0.8 * X[:, k]
    T = np.zeros((n_samples, P))
    for t in range(P):
        par = rng.choice(M, size=min(M, 3), replace=False)
        T[:, t] = X[:, par] @ (rng.uniform(0.15, 0.5, len(par)) * rng.choice([-1, 1], len(par)))
        for u in range(t):
            if rng.random() < trait_density:
                T[:, t] += rng.uniform(-0.6, 0.6) * T[:, u]
        T[:, t] += rng.standard_normal(n_samples)
    C = np.corrcoef(np.concatenate([T, X], axis=1), rowvar=False).astype(np.float32)
    for a, b in collinear:  # merged indices (traits first)
        C[a, b] = C[b, a] = 1.0
    z = np.abs(np.arctanh(np.clip(C.astype(np.float64), -0.999999, 0.999999)))
    A = (z * np.sqrt(n_samples - 3) > 3.89).astype(np.int32)
    np.fill_diagonal(A, 0)
    A[n - 1, 0] = A[0, n - 1] = 1  # write_mm sizes both files by the largest adjacency row index
    sam, scm = {}, {}
    for i in range(n):
        for j in range(n):
            if A[i, j]:
                sam[(np.int64(i + 1), np.int64(j + 1))] = np.int32(1)
            if C[i, j] != 0 and (i != j or rng.random() < 0.8):  # some diagonal entries absent: the reader sets them to 1
                scm[(np.int64(i + 1), np.int64(j + 1))] = C[i, j]
    gmi = {k: 10 * k + 3 for k in range(M)}
    GlobalBdpcResult(sam, scm, gmi, n, P, 14).write_mm(os.path.join(d, "merged_blocks"))


def snapshot(d):
    out = {}
    for fn in sorted(os.listdir(d)):
        with open(os.path.join(d, fn), "rb") as f:
            out[fn] = base64.b64encode(f.read()).decode()
    return out


def case(P, M, seed, n_samples, alpha, prior=None, **kw):
    with tempfile.TemporaryDirectory() as d:
        make_inputs(d, P, M, seed, n_samples, **kw)
        prior_path = None
        if prior is not None:
            prior_path = os.path.join(d, "prior.bin")
            np.asarray(prior, np.int32).tofile(prior_path)
        before = snapshot(d)
        cr = orient_v_structures_merged(d + "/merged_blocks", alpha, n_samples, prior_path)
        cr.to_file(d + "/max_sep_min_pc")
        after = snapshot(d)
        info = {"alpha": alpha, "num_samples": n_samples, "prior": prior is not None, "pairs": len(cr.max_sepsets),
                "pairs_with_minimum": len(cr.min_sepsets), "num_var_after": int(cr.num_var),
                "max_sepset_len": {f"{i},{j}": len(v) for (i, j), v in cr.max_sepsets.items()}}
        print(info["pairs"], "pairs,", info["pairs_with_minimum"], "with a minimum, max level", cr.max_level_maximal_sepsets,
              "ambiguous", cr.ambiguous_triples.shape[0], "ut", cr.get_rfci_relevant_unshielded_triples().shape[0])
        info["max_sepset_len"] = {k: v for k, v in info["max_sepset_len"].items()}
        return {"in": before, "out": {k: v for k, v in after.items() if k not in before}, "info": info}


def main():
    rng = np.random.default_rng(99)
    prior = np.triu((rng.random((6, 6)) < 0.3).astype(np.int32), 1)
    cases = {
        "small": case(6, 30, 1, 20000, 1e-4),
        "traits17": case(17, 60, 2, 50000, 1e-4, trait_density=0.6),
        "collinear": case(8, 40, 3, 20000, 1e-3, collinear=[(10, 11), (20, 30), (31, 32), (2, 15)]),
        "prior": case(6, 30, 1, 20000, 1e-4, prior=prior),
    }
    with open(OUT, "w") as f:
        json.dump(cases, f, indent=0, sort_keys=True)
    print({k: sorted(v["out"]) for k, v in cases.items()})
    print(os.path.getsize(OUT), "bytes")


if __name__ == "__main__":
    main()
