"""world_size-2 gloo tests (CPU) of the multi-GPU host logic: the block partition used by one-process-per-GPU
launches, and the edge-sharded skeleton scheme (ranks run disjoint parts of a level, AND their
adjacency bitmaps, reach the single-rank result) with the oracle standing in for the kernel."""
import os
import socket
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, q):
    sys.path.insert(0, ROOT)
    import torch
    import torch.distributed as dist

    from cigwas_b200 import sched, synth
    from tests import oracle_lib

    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    orc = oracle_lib.load()
    th = orc.hetcor_threshold(1e-4)
    # ---- (1) independent LD blocks: LPT partition, no collective on the data path -----------------
    sizes = [40, 90, 25, 60, 75, 30, 55]
    mine = sched.blocks_of_rank(sizes, rank, world)
    res = {}
    for b in mine:
        Cm = synth.sem_correlation(sizes[b], 3, stream=700 + b, num_samples=20000, causal_per_trait=5)
        n = Cm.shape[0]
        r = orc.hetcor_skeleton(Cm, np.ones((n, n), np.int32), np.full((n, n), 20000.0, np.float32), th, 14,
                                np.zeros(n, np.int32))
        res[b] = int(r["G"].sum())
    gathered = [None] * world
    dist.all_gather_object(gathered, res)
    # ---- (2) one skeleton sharded over the ranks -------------------------------------------------------
    m, p = 120, 4
    Cm = synth.sem_correlation(m, p, stream=731, num_samples=20000, causal_per_trait=6)
    n = m + p
    N = synth.het_sample_sizes(n, m, 20000, stream=732, lo=4000, na_rate=0.0)
    ti = np.zeros(n, np.int32)
    ti[m:] = 1
    G = orc.hetcor_skeleton(Cm, np.ones((n, n), np.int32), N, th, 0, ti)["G"]  # level 0 is replicated
    levels = 0
    for l in range(1, 15):
        Gl = G.copy()
        ran = orc.hetcor_level(Cm, Gl, N, th, l, ti, row_rem=rank, row_mod=world)
        if not ran:
            break
        bits = torch.from_numpy(np.packbits(Gl.astype(np.uint8), axis=1))
        G = np.unpackbits(sched.and_allgather(bits).numpy(), axis=1)[:, :n].astype(np.int32)
        levels = l
    if rank == 0:
        q.put(dict(gathered=gathered, G=G, levels=levels))
    dist.barrier()
    dist.destroy_process_group()


def test_block_partition_properties():
    sys.path.insert(0, ROOT)
    from cigwas_b200 import sched

    rng = np.random.default_rng(1)
    sizes = np.clip(np.round(rng.lognormal(np.log(600), 0.6, 1700)), 50, 3000).astype(int)
    for world in (1, 2, 4, 8):
        parts = sched.lpt_partition(sizes, world)
        flat = sorted(b for part in parts for b in part)
        assert flat == list(range(len(sizes)))
        loads = [sum(float(sizes[b]) ** 2 for b in part) for part in parts]
        assert max(loads) / (sum(loads) / world) < 1.02  # LPT on 1 700 blocks is essentially balanced
        for part in parts:
            assert all(sizes[part[i]] >= sizes[part[i + 1]] for i in range(len(part) - 1))


@pytest.mark.timeout(600)
def test_gloo_world2_blocks_and_sharded_skeleton(oracle):
    import torch.multiprocessing as mp

    from cigwas_b200 import synth

    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for pr in procs:
        pr.start()
    out = q.get(timeout=500)
    for pr in procs:
        pr.join(timeout=60)
        assert pr.exitcode == 0
    # (1) every block processed exactly once, results equal to a serial pass
    sizes = [40, 90, 25, 60, 75, 30, 55]
    merged = {}
    for part in out["gathered"]:
        assert not (set(part) & set(merged))
        merged.update(part)
    assert sorted(merged) == list(range(len(sizes)))
    th = oracle.hetcor_threshold(1e-4)
    for b in (1, 4):
        Cm = synth.sem_correlation(sizes[b], 3, stream=700 + b, num_samples=20000, causal_per_trait=5)
        n = Cm.shape[0]
        r = oracle.hetcor_skeleton(Cm, np.ones((n, n), np.int32), np.full((n, n), 20000.0, np.float32), th, 14,
                                   np.zeros(n, np.int32))
        assert merged[b] == int(r["G"].sum())
    # (2) sharded levels + AND exchange == the single-rank skeleton
    m, p = 120, 4
    Cm = synth.sem_correlation(m, p, stream=731, num_samples=20000, causal_per_trait=6)
    n = m + p
    N = synth.het_sample_sizes(n, m, 20000, stream=732, lo=4000, na_rate=0.0)
    ti = np.zeros(n, np.int32)
    ti[m:] = 1
    full = oracle.hetcor_skeleton(Cm, np.ones((n, n), np.int32), N, th, 14, ti)
    assert (out["G"] == full["G"]).all()
    assert out["levels"] == full["l"] or out["levels"] == min(14, full["l"])
