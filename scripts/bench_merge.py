#!/usr/bin/env python3
"""Time the native merge-block-outputs (mps merge) on synthetic whole-genome block outputs (C4-like: 1 700 blocks,
17 traits, a few dozen retained markers per block), and the reference's Python next to it when
cusk_postprocessing is importable (build container only; pass its root with --reference).  Checks that both write
the same bytes.  Host-only: no GPU involved."""
import argparse
import filecmp
import json
import os
import subprocess
import sys
import tempfile
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def make_outputs(d, blocks, traits, mean_markers, seed):
    rng = np.random.default_rng(seed)
    first, lines, nbytes = 0, [], 0
    for _ in range(blocks):
        blen = int(rng.integers(200, 3000))
        m = int(np.clip(rng.poisson(mean_markers), 1, blen))
        n = m + traits
        a = np.triu((rng.random((n, n)) < 4.0 / n).astype(np.int32), 1)
        a = a + a.T
        c = np.triu(rng.uniform(-1, 1, (n, n)).astype(np.float32), 1)
        c = c + c.T
        np.fill_diagonal(c, 1.0)
        stem = os.path.join(d, f"1_{first}_{first + blen - 1}")
        open(stem + ".mdim", "w").write(f"{n}\t{traits}\t14\n")
        a.tofile(stem + ".adj")
        c.tofile(stem + ".corr")
        np.concatenate([np.sort(rng.choice(blen, m, replace=False)), blen + np.arange(traits)]).astype(np.int32).tofile(stem + ".ixs")
        nbytes += 8 * n * n + 4 * n
        lines.append(f"1\t{first}\t{first + blen - 1}\n")
        first += blen
    open(os.path.join(d, "all.blocks"), "w").writelines(lines)
    return nbytes


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--blocks", type=int, default=1700)
    ap.add_argument("--traits", type=int, default=17)
    ap.add_argument("--mean-markers", type=float, default=40.0)
    ap.add_argument("--reference", default="/root/reference")
    args = ap.parse_args()
    import cigwas_b200

    with tempfile.TemporaryDirectory() as d:
        nbytes = make_outputs(d, args.blocks, args.traits, args.mean_markers, 20261104)
        t0 = time.perf_counter()
        subprocess.run([cigwas_b200.mps_path(), "merge", os.path.join(d, "all.blocks"), d], check=True, stdout=subprocess.DEVNULL)
        t_native = time.perf_counter() - t0
        out = {"blocks": args.blocks, "traits": args.traits, "input_bytes": nbytes, "native_s": t_native,
               "native_MB_per_s": nbytes / t_native / 1e6,
               "scm_entries": int(open(os.path.join(d, "merged_blocks_scm.mtx")).readlines()[1].split()[2])}
        if os.path.isdir(os.path.join(args.reference, "cusk_postprocessing")):
            sys.path.insert(0, args.reference)
            from cusk_postprocessing.merge_blocks import merge_block_outputs

            t0 = time.perf_counter()
            merge_block_outputs(os.path.join(d, "all.blocks"), d + "/").write_mm(d + "/ref_blocks")
            out["reference_python_s"] = time.perf_counter() - t0
            out["identical_bytes"] = all(filecmp.cmp(f"{d}/merged_blocks{s}", f"{d}/ref_blocks{s}", shallow=False)
                                         for s in ("_sam.mtx", "_scm.mtx", ".mdim", ".ixs"))
            out["speedup"] = out["reference_python_s"] / t_native
        print(json.dumps(out))


if __name__ == "__main__":
    main()
