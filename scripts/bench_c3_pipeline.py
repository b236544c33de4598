#!/usr/bin/env python
"""BASELINE.json configs[2] (C3) as a PIPELINE: `mps prep` + `mps cusk` on individual-level data from a packed BED file,
458 000 samples x 10 000 markers + 17 traits in one block (cli.cpp:432-678: marker x trait screen -> all correlations ->
square matrix -> two-stage skeleton -> .mdim/.ixs/.adj/.corr/.sep), through the `mps` binary as `ci-gwas.py cusk` calls it.

    python scripts/bench_c3_pipeline.py [--n 458000] [--m 10000] [--steps 2] [--check-pairs 200]

One JSON line: wall seconds of prep and of cusk (files in, files out, page cache warm), the block line of the scheduler
(load / device / write split), BED GB/s over the whole pipeline, and an output check at this size: the run with level 0
fused into the correlation epilogue against the run with level 0 as its own pass (files must be byte-identical), and
sampled entries of the written .corr against the CPU oracle's pair tables (1e-6).
"""
import argparse
import filecmp
import json
import os
import shutil
import subprocess
import sys
import tempfile
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
MPS = os.path.join(ROOT, "ci-gwas_b200", "mps")


def run(cmd, env=None):
    t0 = time.perf_counter()
    r = subprocess.run([str(c) for c in cmd], capture_output=True, text=True, env=dict(os.environ, **(env or {})))
    dt = time.perf_counter() - t0
    if r.returncode != 0:
        raise SystemExit(f"{cmd[:2]} failed ({r.returncode}): {(r.stderr or r.stdout)[-1500:]}")
    return dt, r.stdout


def main():
    import bench

    ap = argparse.ArgumentParser()
    ap.add_argument("--n", type=int, default=458000)
    ap.add_argument("--m", type=int, default=10000)
    ap.add_argument("--p", type=int, default=17)
    ap.add_argument("--steps", type=int, default=2)
    ap.add_argument("--check-pairs", type=int, default=200)
    a = ap.parse_args()
    n, m, p = a.n, a.m, a.p
    base = tempfile.mkdtemp(prefix="c3_", dir="/dev/shm" if os.path.isdir("/dev/shm") else None)
    stem = os.path.join(base, "c3")
    t0 = time.perf_counter()
    bed = bench.fast_bed(n, m, 3)
    with open(stem + ".bed", "wb") as f:
        f.write(bytes([0x6C, 0x1B, 0x01]))
        f.write(bed.tobytes())
    with open(stem + ".bim", "w") as f:
        f.write("".join(f"1\trs{i}\t0\t{i + 1}\tA\tC\n" for i in range(m)))
    with open(stem + ".fam", "w") as f:
        f.write("".join(f"f{i} i{i} 0 0 0 -9\n" for i in range(n)))
    rng = np.random.Generator(np.random.PCG64([20261104, 99]))
    ph = rng.standard_normal((n, p), dtype=np.float32)
    # a few real marker -> trait effects so that the block passes the screen on signal, not on chance
    code = (bed[:8, :].reshape(8, -1, 1) >> np.array([0, 2, 4, 6], np.uint8)) & 3
    g = np.array([2, 0, 1, 0], np.float32)[code.reshape(8, -1)[:, :n]]
    for t in range(p):
        gt = g[t % 8]
        ph[:, t] += 0.05 * (gt - gt.mean()) / gt.std()
    ph = (ph - ph.mean(0)) / ph.std(0)
    cols = [np.char.mod("%.6g", ph[:, t]) for t in range(p)]
    for t in range(3):
        cols[t][rng.random(n) < 0.01] = "NA"
    ids = np.char.add("f", np.arange(n).astype(str))
    rows = ids
    for c in [np.char.add("i", np.arange(n).astype(str))] + cols:
        rows = np.char.add(np.char.add(rows, " "), c)
    with open(os.path.join(base, "traits.phen"), "w") as f:
        f.write("FID IID " + " ".join(f"T{t + 1}" for t in range(p)) + "\n")
        f.write("\n".join(rows.tolist()) + "\n")
    open(os.path.join(base, "c3.blocks"), "w").write(f"1\t0\t{m - 1}\n")
    gen_s = time.perf_counter() - t0
    env = {"CUSK_DEVICES": "0", "CUSK_NO_RESUME": "1", "CUSK_SLOTS_PER_GPU": "1"}
    prep_s, _ = run([MPS, "prep", stem], env)
    args = [os.path.join(base, "traits.phen"), stem, os.path.join(base, "c3.blocks"), 1e-4, 3, 14, 1]
    walls, block_line = [], None
    for i in range(a.steps + 1):  # first run warms the page cache and the device
        out = os.path.join(base, f"out{i}")
        os.makedirs(out)
        dt, so = run([MPS, "cusk-all"] + args + [out], env)
        if i:
            walls.append(dt)
        for ln in so.splitlines():
            if ln.startswith('{"block"'):
                block_line = json.loads(ln)
    # level 0 as its own pass: same files
    out_nf = os.path.join(base, "out_nofuse")
    os.makedirs(out_nf)
    run([MPS, "cusk-all"] + args + [out_nf], dict(env, CUSK_FUSE_LEVEL0="0"))
    last = os.path.join(base, f"out{a.steps}")
    sfiles = [f for f in sorted(os.listdir(last))]
    same = sorted(os.listdir(out_nf)) == sfiles and all(filecmp.cmp(os.path.join(last, f), os.path.join(out_nf, f), False) for f in sfiles)
    # sampled entries of the written correlations against the oracle's exact tables
    check = {"files": sfiles, "fused_equals_unfused": bool(same)}
    s0 = os.path.join(last, f"1_0_{m - 1}")
    if os.path.exists(s0 + ".mdim") and a.check_pairs > 0:
        from tests import oracle_lib

        orc = oracle_lib.load()
        nv, npn, _ = [int(x) for x in open(s0 + ".mdim").read().split()]
        ixs = np.fromfile(s0 + ".ixs", np.int32)
        corr = np.fromfile(s0 + ".corr", np.float32).reshape(nv, nv)
        adj = np.fromfile(s0 + ".adj", np.int32).reshape(nv, nv)
        mk = nv - npn
        rs = np.random.default_rng(5)
        bad, done = 0, 0
        for _ in range(a.check_pairs):
            if mk < 2:
                break
            i, j = rs.choice(mk, size=2, replace=False)
            t = orc.pair_table(bed, n, int(ixs[i]), int(ixs[j]))
            e = orc.npn_from_table(t)
            gval = float(corr[i, j])
            ok = (np.isnan(e) and np.isnan(gval)) or abs(gval - e) <= 1e-6 * abs(e) + 2e-7
            bad += 0 if ok else 1
            done += 1
        check.update({"num_var_out": nv, "markers_out": mk, "edges_out": int(adj.sum() // 2),
                      "corr_entries_checked_vs_oracle": done, "corr_entries_outside_1e-6": bad})
    wall = float(np.mean(walls))
    bed_gb = m * ((n + 3) // 4) / 1e9
    print(json.dumps({
        "metric": "cusk_pipeline_seconds_per_block", "workload": f"`mps prep` + `mps cusk-all` (one block): {n} samples x {m} "
        f"markers + {p} traits from a packed .bed (BASELINE.json configs[2]), alpha 1e-4, max levels 3 / 14, depth 1",
        "generation_s": round(gen_s, 1), "prep_s": round(prep_s, 3), "cusk_wall_s": [round(w, 3) for w in walls],
        "cusk_wall_s_mean": wall, "bed_gb": bed_gb, "bed_gbps_whole_pipeline": bed_gb / wall,
        "block": block_line, "check": check}), flush=True)
    shutil.rmtree(base, ignore_errors=True)


if __name__ == "__main__":
    main()
