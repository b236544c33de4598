#!/usr/bin/env python
"""BASELINE.json configs[3] (C4): whole-genome cuskss -- ~1 700 synthetic LD blocks through the block scheduler
(`mps cuskss-all`: LPT queue, CUSK_SLOTS_PER_GPU host threads per GPU so that file IO / assembly / copies of one block
overlap the device phase of another, no collective between GPUs).  Also reachable as `bench.py --workload c4`.

    python scripts/bench_c4.py [--gpus N] [--blocks 1700] [--steps 2] [--warmup 1] [--levels 14,0] [--verify 20] [--ref-sample 50]

Under torchrun (N ranks) rank 0 drives all N GPUs from the one `mps` process, as a user of the CLI would; the other
ranks only wait.  One JSON line: value = CI tests per second of wall-clock over a whole `cuskss-all` run (files in, files
out; page cache warm), the scheduler's own summary (per-GPU device-phase share, load / device / write split), the same
blocks through per-block `mps cuskss` processes (the reference's model: one process per block, README.md:62) for a
byte comparison on a sample, and the reference's own hetcor_skeleton kernels (oracle/_ref) on a sample of blocks.
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
ALPHA, NS = 1e-4, 458000


def mps_path():
    return os.path.join(ROOT, "ci-gwas_b200", "mps")


def run_all(gdir, out, gpus, levels, env_extra=None):
    env = dict(os.environ, CUSK_DEVICES=",".join(str(i) for i in range(gpus)), CUSK_NO_RESUME="1",
               OMP_NUM_THREADS=str(min(32, os.cpu_count() or 1)))
    env.update(env_extra or {})
    cmd = [mps_path(), "cuskss-all", os.path.join(gdir, "mxm", "{}.bin"), os.path.join(gdir, "mxp.txt"), "NULL",
           os.path.join(gdir, "pxp.txt"), "NULL", "NULL", os.path.join(gdir, "genome.blocks"), str(ALPHA), str(levels[0]),
           str(levels[1]), "1", str(NS), out]
    t0 = time.perf_counter()
    r = subprocess.run(cmd, capture_output=True, text=True, env=env)
    wall = time.perf_counter() - t0
    if r.returncode != 0:
        raise SystemExit(f"mps cuskss-all failed ({r.returncode}): {r.stderr[-2000:]}")
    summary, blocks = None, []
    for ln in r.stdout.splitlines():
        if ln.startswith("{"):
            d = json.loads(ln)
            if d.get("summary"):
                summary = d
            else:
                blocks.append(d)
    return wall, summary, blocks


def block_matrix(gdir, info, b, mxp_rows, pxp):
    """the square correlation matrix `mps cuskss` assembles for block b (cli.cpp:89-173)"""
    m = int(info["sizes"][b])
    tril = np.fromfile(os.path.join(gdir, "mxm", info["stems"][b] + ".bin"), np.float32)
    p = pxp.shape[0]
    C = np.ones((m + p, m + p), np.float32)
    il = np.tril_indices(m)
    C[:m, :m][il] = tril
    C[:m, :m].T[il] = tril
    C[:m, m:] = mxp_rows
    C[m:, :m] = mxp_rows.T
    C[m:, m:] = pxp
    return C


def main(argv=None):
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--blocks", type=int, default=1700)
    ap.add_argument("--steps", type=int, default=2)
    ap.add_argument("--warmup", type=int, default=1)
    ap.add_argument("--levels", default="14,0", help="max-level-one,max-level-two (0: single stage)")
    ap.add_argument("--verify", type=int, default=20, help="blocks re-run as single `mps cuskss` processes and compared")
    ap.add_argument("--ref-sample", type=int, default=50, help="blocks run through the reference's own kernels")
    ap.add_argument("--slots", type=int, default=0, help="CUSK_SLOTS_PER_GPU (0: library default)")
    ap.add_argument("--keep", default="", help="directory to generate into and keep (default: a temporary one)")
    args, _ = ap.parse_known_args(argv)
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    gpus = max(args.gpus, world)
    if world > 1:
        import torch.distributed as dist

        dist.init_process_group("gloo")
    if rank == 0:
        line = run(args, gpus)
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


def run(args, gpus):
    from cigwas_b200 import synth

    levels = [int(x) for x in args.levels.split(",")]
    base = args.keep or tempfile.mkdtemp(prefix="c4_", dir="/dev/shm" if os.path.isdir("/dev/shm") else None)
    gdir = os.path.join(base, "genome")
    t0 = time.perf_counter()
    info = synth.c4_genome(gdir, nblocks=args.blocks, num_samples=NS)
    gen_s = time.perf_counter() - t0
    env_extra = {"CUSK_SLOTS_PER_GPU": str(args.slots)} if args.slots > 0 else {}
    outs = []
    for i in range(args.warmup):
        o = os.path.join(base, f"warm{i}")
        os.makedirs(o)
        run_all(gdir, o, gpus, levels, env_extra)
        shutil.rmtree(o)
    walls, summaries = [], []
    for i in range(args.steps):
        o = os.path.join(base, f"out{i}")
        os.makedirs(o)
        wall, summary, blocks = run_all(gdir, o, gpus, levels, env_extra)
        walls.append(wall)
        summaries.append(summary)
        outs.append(o)
    tests = int(summaries[-1]["ci_tests"])
    wall = float(np.mean(walls))
    line = {
        "metric": "ci_tests_per_s", "value": tests / wall, "unit": "tests/s", "n_gpus": gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": 1e3 * wall, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
        "dtype": "f64", "data": "synthetic",
        "config": {"workload": f"whole-genome cuskss: {args.blocks} synthetic LD blocks ({info['markers']} markers, sizes "
                               "clip(LogNormal(ln 600, 0.6), 50, 3000)) + 17 traits through `mps cuskss-all`, max levels "
                               f"{levels[0]}/{levels[1]}, alpha 1e-4, 458000 samples (BASELINE.json configs[3]); files in, files out",
                   "generator": "cigwas_b200.synth.c4_genome (C2 generator per block, own stream; shared traits)",
                   "generation_s": round(gen_s, 1), "l2": "inputs come from files (page cache warm) every step"},
        "wall_s_per_run": [round(w, 3) for w in walls],
        "scheduler": summaries[-1],
        "scheduler_seconds": round(float(summaries[-1]["seconds"]), 3),
        "processing_s": summaries[-1].get("processing_s"),
        "startup_s": round(wall - float(summaries[-1].get("processing_s", 0.0)), 3),
        "tests_per_s_processing": tests / float(summaries[-1]["processing_s"]) if summaries[-1].get("processing_s") else None,
        "note": "wall = the whole `mps cuskss-all` process (CUDA initialisation of every GPU, the shared marker x trait table, "
                "all blocks, teardown); processing_s = from the moment every worker has its context and the shared tables "
                "are loaded to the last block written: the part that block-parallel scheduling divides over the GPUs",
    }
    line["e2e"] = {"value": tests / wall, "unit": "tests/s", "ms_per_step": 1e3 * wall,
                   "note": "the step IS end to end: one `mps cuskss-all` process, input files to output files",
                   "h2d_bytes_per_step": None, "d2h_bytes_per_step": None}
    # ---- byte comparison with one `mps cuskss` process per block (the reference's execution model) -----------------
    rs = np.random.default_rng(7)
    if args.verify > 0:
        sample = sorted(rs.choice(args.blocks, size=min(args.verify, args.blocks), replace=False).tolist())
        vdir = os.path.join(base, "verify")
        os.makedirs(vdir)
        t0 = time.perf_counter()
        same, compared = True, 0
        for b in sample:
            stem = info["stems"][b]
            cmd = [mps_path(), "cuskss", os.path.join(gdir, "mxm", stem + ".bin"), os.path.join(gdir, "mxp.txt"), "NULL",
                   os.path.join(gdir, "pxp.txt"), "NULL", "NULL", str(b), os.path.join(gdir, "genome.blocks"), "NULL", str(ALPHA),
                   str(levels[0]), str(levels[1]), "1", str(NS), vdir]
            r = subprocess.run(cmd, capture_output=True, text=True, env=dict(os.environ, CUSK_DEVICES="0"))
            if r.returncode != 0:
                raise SystemExit(f"mps cuskss failed on block {b}: {r.stderr[-1000:]}")
            for sfx in (".mdim", ".ixs", ".adj", ".corr"):
                compared += 1
                same = same and filecmp.cmp(os.path.join(vdir, stem + sfx), os.path.join(outs[-1], stem + sfx), shallow=False)
        per_block_s = (time.perf_counter() - t0) / len(sample)
        line["per_block_processes"] = {"blocks": len(sample), "files_compared": compared, "byte_identical": bool(same),
                                       "seconds_per_block_process": round(per_block_s, 3),
                                       "extrapolated_s_all_blocks_1gpu": round(per_block_s * args.blocks, 1)}
    # ---- the reference's own kernels on a sample of blocks (same box) -------------------------------------------
    if args.ref_sample > 0:
        try:
            from tests import ref_lib

            if ref_lib.available():
                import bench as B
                import cigwas_b200

                ref = ref_lib.RefLib()
                ctx = cigwas_b200.Context(0)
                th = cigwas_b200.hetcor_threshold(ALPHA)
                pxp = np.loadtxt(os.path.join(gdir, "pxp.txt"), skiprows=1, usecols=range(1, 18), dtype=np.float32)
                mxp_all = np.loadtxt(os.path.join(gdir, "mxp.txt"), skiprows=1, usecols=range(3, 20), dtype=np.float32)
                starts = np.concatenate([[0], np.cumsum(info["sizes"])[:-1]])
                sample = sorted(rs.choice(args.blocks, size=min(args.ref_sample, args.blocks), replace=False).tolist())
                ref_s, our_s, diff, ref_tests = 0.0, 0.0, 0, 0
                for b in sample:
                    m = int(info["sizes"][b])
                    C = block_matrix(gdir, info, b, mxp_all[starts[b]:starts[b] + m], pxp)
                    n = C.shape[0]
                    N = np.full((n, n), float(NS), np.float32)
                    ti = np.zeros(n, np.int32)
                    ti[m:] = 1
                    G1 = np.ones((n, n), np.int32)
                    with B.quiet_stdout():
                        t0 = time.perf_counter()
                        rr = ref.hetcor_skeleton(C, G1, N, th, levels[0], ti)
                        ref_s += time.perf_counter() - t0
                    t0 = time.perf_counter()
                    ro = ctx.hetcor_skeleton(C, G1, N, th, levels[0], ti)
                    our_s += time.perf_counter() - t0
                    diff += int((rr["G"] != ro["G"]).sum())
                    ref_tests += sum(ro["stats"]["tests_executed"])
                ctx.close()
                line["ref_gpu"] = {"what": "the reference's own hetcor_skeleton (unmodified sources, sm_100 build) on a sample of "
                                           "the blocks, first stage, host buffers in/out, wall clock, one GPU",
                                   "blocks": len(sample), "ref_seconds": round(ref_s, 3),
                                   "ours_same_calls_seconds": round(our_s, 3),
                                   "adjacency_entries_differing": diff,
                                   "extrapolated_ref_s_all_blocks": round(ref_s * args.blocks / len(sample), 1)}
        except Exception as e:  # noqa: BLE001
            line["ref_gpu"] = {"unavailable": repr(e)[:300]}
    if not args.keep:
        shutil.rmtree(base, ignore_errors=True)
    return line


if __name__ == "__main__":
    main()
