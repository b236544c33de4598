#!/usr/bin/env python3
"""Golden input/output pairs for `merge-block-outputs` and the cuskss-merged reformatting step, made by RUNNING the
reference's own Python (cusk_postprocessing/merge_blocks.py: merge_block_outputs :361-403, write_mm :300-325,
reformat_cuskss_merged_output :406-425) on small synthetic block outputs.

Run ONCE in the build container (where /root/reference is mounted); the result, tests/golden/merge_golden.json,
is committed so that no test needs /root/reference at run time.  Every file is stored base64-encoded.

Cases (3 traits unless stated):
  four_blocks     4 blocks, the second one without outputs; trait x trait links that differ between blocks
                  (intersection for traits 1..P-1, union for trait P: the reference's range(num_p) loop), trait x trait
                  correlations that change and appear between blocks, tiny / negative / NaN / integer-valued correlations
  first_missing   the first block has no outputs (the reference then never adds links among traits 1..P-1)
  dense_random    6 blocks of random size with 5 traits, seeded
  reformat        cuskss_merged.{mdim,ixs,adj,corr} + merged_blocks.ixs -> cuskss_merged_{sam,scm}.mtx, .mdim, .ixs
"""
import base64
import json
import os
import sys
import tempfile

import numpy as np

sys.path.insert(0, os.environ.get("CIGWAS_REFERENCE", "/root/reference"))
from cusk_postprocessing.merge_blocks import merge_block_outputs, reformat_cuskss_merged_output  # noqa: E402

OUT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "merge_golden.json")


def write_block(d, stem, num_m, num_p, adj, corr, rel, max_level=14):
    n = num_m + num_p
    with open(os.path.join(d, stem + ".mdim"), "w") as f:
        f.write(f"{n}\t{num_p}\t{max_level}\n")
    np.asarray(adj, np.int32).reshape(n, n).tofile(os.path.join(d, stem + ".adj"))
    np.asarray(corr, np.float32).reshape(n, n).tofile(os.path.join(d, stem + ".corr"))
    np.asarray(rel, np.int32).tofile(os.path.join(d, stem + ".ixs"))


def random_block(rng, num_m, num_p, block_len, density=0.4):
    n = num_m + num_p
    a = (rng.random((n, n)) < density).astype(np.int32)
    a = np.triu(a, 1)
    a = a + a.T
    c = rng.uniform(-1, 1, (n, n)).astype(np.float32)
    c = np.triu(c, 1)
    c = c + c.T
    np.fill_diagonal(c, 1.0)
    c[rng.random((n, n)) < 0.1] = 0.0
    rel = np.concatenate([np.sort(rng.choice(block_len, num_m, replace=False)), block_len + np.arange(num_p)])
    return a, c, rel


def snapshot(d):
    out = {}
    for fn in sorted(os.listdir(d)):
        with open(os.path.join(d, fn), "rb") as f:
            out[fn] = base64.b64encode(f.read()).decode()
    return out


def merge_case(blocks, make):
    """blocks: [(chr, first, last)]; make(d) writes the block outputs; returns {"in": files, "out": files}"""
    with tempfile.TemporaryDirectory() as d:
        with open(os.path.join(d, "all.blocks"), "w") as f:
            for b in blocks:
                f.write("\t".join(str(x) for x in b) + "\n")
        make(d)
        before = snapshot(d)
        merge_block_outputs(os.path.join(d, "all.blocks"), d + "/").write_mm(d + "/merged_blocks")
        after = snapshot(d)
        return {"in": before, "out": {k: v for k, v in after.items() if k not in before}}


def four_blocks(d):
    P = 3
    # block 0: 2 markers
    a = np.zeros((5, 5), np.int32)
    for i, j in [(0, 2), (1, 3), (0, 1), (2, 3), (2, 4), (3, 4)]:  # m0-t1, m1-t2, m0-m1, t1-t2, t1-t3, t2-t3
        a[i, j] = a[j, i] = 1
    c = np.array([[1, 0.5, 1e-5, -0.25, 0], [0.5, 1, 0.299969, 1.5e-5, 1e-4], [1e-5, 0.299969, 1, 0.125, np.nan],
                  [-0.25, 1.5e-5, 0.125, 1, 2.0], [0, 1e-4, np.nan, 2.0, 1]], np.float32)
    write_block(d, "1_0_9", 2, P, a, c, [3, 7, 10, 11, 12])
    # block 1 (1_10_29): no outputs
    # block 2: 1 marker; t1-t2 absent (deleted), t1-t3 absent (kept: trait P is never intersected), t2-t3 present
    a = np.zeros((4, 4), np.int32)
    for i, j in [(0, 3), (2, 3)]:  # m0-t3, t2-t3
        a[i, j] = a[j, i] = 1
    c = np.array([[1, 0.75, 0, -1e-7], [0.75, 1, 0.25, 0.0625], [0, 0.25, 1, 123456.7], [-1e-7, 0.0625, 123456.7, 1]],
                 np.float32)
    write_block(d, "1_30_34", 1, P, a, c, [2, 5, 6, 7])
    # block 3: 3 markers on another chromosome; new trait x trait links t1-t2 (not re-added) and t1-t3 (updated)
    a = np.zeros((6, 6), np.int32)
    for i, j in [(0, 1), (1, 2), (2, 5), (3, 4), (3, 5), (0, 3)]:
        a[i, j] = a[j, i] = 1
    rng = np.random.default_rng(7)
    c = rng.uniform(-1, 1, (6, 6)).astype(np.float32)
    c = (c + c.T) / 2
    np.fill_diagonal(c, 1.0)
    c[3, 4] = c[4, 3] = 0.0
    write_block(d, "2_0_19", 3, P, a, c, [0, 8, 19, 20, 21, 22])


def first_missing(d):
    rng = np.random.default_rng(11)
    for stem, m, blen in [("1_5_24", 3, 20), ("1_25_30", 2, 6)]:
        a, c, rel = random_block(rng, m, 3, blen, 0.6)
        write_block(d, stem, m, 3, a, c, rel)


def dense_random(d):
    rng = np.random.default_rng(20261104)
    first = 0
    for k in range(6):
        blen = int(rng.integers(8, 30))
        m = int(rng.integers(1, 7))
        a, c, rel = random_block(rng, m, 5, blen, 0.5)
        write_block(d, f"3_{first}_{first + blen - 1}", m, 5, a, c, rel, max_level=6)
        first += blen


def dense_blocks_from_dir(make):
    """run `make` once to learn the stems it writes"""
    with tempfile.TemporaryDirectory() as d:
        make(d)
        stems = sorted((fn[:-5] for fn in os.listdir(d) if fn.endswith(".mdim")), key=lambda s: int(s.split("_")[1]))
        return [tuple(s.split("_")) for s in stems]


def reformat_case():
    with tempfile.TemporaryDirectory() as d:
        rng = np.random.default_rng(5)
        m, p = 4, 3
        a, c, _ = random_block(rng, m, p, 10, 0.5)
        write_block(d, "cuskss_merged", m, p, a, c, [1, 2, 5, 6, 7, 8, 9])  # selection-relative indices, then traits
        np.array([4, 9, 12, 100, 101, 250, 251], np.int32).tofile(os.path.join(d, "merged_blocks.ixs"))
        before = snapshot(d)
        reformat_cuskss_merged_output(cusk_dir=d).write_mm(basepath=f"{d}/cuskss_merged")
        after = snapshot(d)
        return {"in": before, "out": {k: v for k, v in after.items() if before.get(k) != v}}


def main():
    cases = {
        "four_blocks": merge_case([(1, 0, 9), (1, 10, 29), (1, 30, 34), (2, 0, 19)], four_blocks),
        "first_missing": merge_case([(1, 0, 4), (1, 5, 24), (1, 25, 30)], first_missing),
        "dense_random": merge_case(dense_blocks_from_dir(dense_random), dense_random),
        "reformat": reformat_case(),
    }
    with open(OUT, "w") as f:
        json.dump(cases, f, indent=0, sort_keys=True)
    for k, v in cases.items():
        print(k, "inputs", len(v["in"]), "outputs", sorted(v["out"]))


if __name__ == "__main__":
    main()
