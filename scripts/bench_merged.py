#!/usr/bin/env python
"""BASELINE.json configs[4]: cuskss-merged on ~20k selected markers + 17 traits (heterologous sample sizes),
the level-l tests of the ONE skeleton sharded over the GPUs of a box, NCCL all-gather of the removed-edge flags
(one byte per CSR slot) between levels.

    python scripts/bench_merged.py                      # 1 GPU
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P \
        scripts/bench_merged.py [--loci 400] [--steps 3] [--warmup 1]

Strong scaling: the problem is fixed, time = max over ranks (CUDA events on each rank's stream), the adjacency
must be identical on every rank and for every N (sha1 of the bitmap is printed).  One JSON line from rank 0.
"""
import argparse
import hashlib
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--loci", type=int, default=400)
    ap.add_argument("--per-locus", type=int, default=50)
    ap.add_argument("--traits", type=int, default=17)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=1)
    ap.add_argument("--maxlevel", type=int, default=14)
    ap.add_argument("--uniform", action="store_true", help="uniform sample size (plain cuskss-merged, no het)")
    ap.add_argument("--causal", type=int, default=12, help="causal markers per trait (30+: trait rows keep ~30 neighbours "
                    "at every level, the hub-heavy regime)")
    ap.add_argument("--beta-lo", type=float, default=0.01)
    ap.add_argument("--beta-hi", type=float, default=0.05)
    ap.add_argument("--tag", default="")
    args = ap.parse_args()
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))

    import torch
    import torch.distributed as dist

    import cigwas_b200
    from cigwas_b200 import synth

    torch.cuda.set_device(local_rank)
    if world > 1:
        saved = os.dup(1)  # NCCL prints its version banner on stdout; keep stdout to the one JSON line
        os.dup2(2, 1)
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
        dist.barrier()
        sys.stdout.flush()
        os.dup2(saved, 1)
        os.close(saved)
    ctx = cigwas_b200.Context(local_rank)
    if world > 1:
        ids = [cigwas_b200.comm_unique_id() if rank == 0 else None]
        dist.broadcast_object_list(ids, src=0)
        ctx.attach_comm(rank, world, ids[0])

    m = args.loci * args.per_locus
    p = args.traits
    n = m + p
    ns = 458000
    Cm = synth.merged_correlation(args.loci, args.per_locus, p, stream=5, num_samples=ns, causal_per_trait=args.causal,
                                  beta_lo=args.beta_lo, beta_hi=args.beta_hi)
    dC = torch.from_numpy(Cm).cuda()
    del Cm
    if args.uniform:
        dN = None
    else:
        N = synth.het_sample_sizes(n, m, ns, stream=5, lo=50000, na_rate=0.005)
        dN = torch.from_numpy(N).cuda()
        del N
    ti = np.zeros(n, np.int32)
    ti[m:] = 1  # cuskss default: markers time 0, traits time 1 (cli.cpp:37-42, :219)
    dT = torch.from_numpy(ti).cuda()
    rw = ctx.row_words(n)
    dbits = torch.zeros((n, rw), dtype=torch.int32, device="cuda")
    th = cigwas_b200.hetcor_threshold(1e-4)
    stream = torch.cuda.ExternalStream(ctx.stream, device=local_rank)
    torch.cuda.synchronize()

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def step():
        return ctx.hetcor_skeleton_dev(dC.data_ptr(), n, dbits.data_ptr(), True, dN.data_ptr() if dN is not None else None,
                                       float(ns), th, args.maxlevel, dT.data_ptr())

    for _ in range(args.warmup):
        step()
    ms, tests, xms, xbytes, st = 0.0, 0, 0.0, 0, None
    barrier()
    for _ in range(args.steps):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream)
        _, st = step()
        e1.record(stream)
        e1.synchronize()
        ms += e0.elapsed_time(e1)
        tests += sum(st["tests_executed"][1:])
        if world > 1:
            cs = ctx.comm_stats()
            xms += cs["exchange_ms"]
            xbytes += cs["exchange_bytes"]
    barrier()
    ctx.set_level_timing(True)  # level times from one more, instrumented skeleton
    _, st = step()
    ctx.set_level_timing(False)
    barrier()
    digest = hashlib.sha1(dbits.cpu().numpy().tobytes()).hexdigest()
    edges = st["edges_after"][max(0, st["levels_run"] - 1)]
    if world > 1:
        t = torch.tensor([ms, xms], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms, xms = t.tolist()
        c = torch.tensor([tests], dtype=torch.int64, device="cuda")
        dist.all_reduce(c, op=dist.ReduceOp.SUM)
        tests = int(c.item())
        ds = [None] * world
        dist.all_gather_object(ds, digest)
        same = all(d == ds[0] for d in ds)
    else:
        same = True
    if rank == 0:
        print(json.dumps({
            "metric": "ci_tests_per_s", "tag": args.tag, "generator": {"loci": args.loci, "per_locus": args.per_locus,
            "causal_per_trait": args.causal, "beta": [args.beta_lo, args.beta_hi], "stream": 5},
            "knobs": {k: v for k, v in os.environ.items() if k.startswith("CUSK_")},
            "tests_executed_per_level": st["tests_executed"], "tests_canonical_per_level": st["tests_canonical"],
            "workload": f"cuskss-merged {m} markers + {p} traits (n_var {n}), "
            f"{'uniform N' if args.uniform else 'heterologous N (0.5% NA)'}, max level {args.maxlevel}, alpha 1e-4; level-l "
            "work items dealt round-robin over the GPUs, ncclAllGather of one removed-flag byte per CSR slot per level",
            "n_gpus": world, "scaling": "strong", "steps": args.steps, "ms_per_skeleton": ms / args.steps,
            "value": tests / (ms * 1e-3), "unit": "tests/s", "tests_per_skeleton": tests // args.steps,
            "levels_run": st["levels_run"], "last_level": st["last_level"], "max_degree": st["max_degree"],
            "edges": edges, "level_ms_rank0": [round(x, 3) for x in st["level_ms"]],
            "exchange_ms_per_skeleton": xms / args.steps, "exchange_bytes_per_skeleton": xbytes // max(1, args.steps),
            "exchange_gbps": (xbytes / 1e9) / (xms * 1e-3) if xms > 0 else None,
            "adjacency_sha1": digest, "adjacency_identical_on_all_ranks": same}), flush=True)
    ctx.close()
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
