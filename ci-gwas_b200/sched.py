"""Multi-GPU plumbing for one-process-per-GPU launches (torchrun): which LD blocks a rank owns, and
the exchange step of the edge-sharded skeleton expressed on torch.distributed (used by the CPU/gloo
tests of the scheme; on GPUs the library does the same exchange with NCCL itself, csrc/sharded.cu).

LD blocks are independent units: no collective is needed between ranks, only a
deterministic, balanced assignment.  Cost model of a block: m^2 (level 0 and the correlation Gram
are quadratic in the block size).
"""
import numpy as np


def lpt_partition(sizes, world):
    """Longest-processing-time-first: blocks sorted by descending cost go to the least loaded rank.
    Returns a list of `world` lists of block indices (each in processing order)."""
    sizes = np.asarray(sizes, np.int64)
    order = np.argsort(-sizes, kind="stable")
    load = np.zeros(world, np.float64)
    parts = [[] for _ in range(world)]
    for b in order:
        r = int(np.argmin(load))
        parts[r].append(int(b))
        load[r] += float(sizes[b]) ** 2
    return parts


def blocks_of_rank(sizes, rank, world):
    return lpt_partition(sizes, world)[rank]


def and_allgather(bits, group=None):
    """Exchange step of a sharded level: every rank only clears bits of its copy of the adjacency
    bitmap, so the common graph is the AND over ranks.  `bits`: integer torch tensor (any device the
    process group supports); returns the combined tensor."""
    import torch
    import torch.distributed as dist

    world = dist.get_world_size(group)
    if world == 1:
        return bits
    parts = [torch.empty_like(bits) for _ in range(world)]
    dist.all_gather(parts, bits.contiguous(), group=group)
    out = parts[0]
    for t in parts[1:]:
        out = torch.bitwise_and(out, t)
    return out
