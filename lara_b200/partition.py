"""Static length-balanced partition of the all-pairs workload over ranks (one rank per GPU).

The reference keeps one longest-first queue (CompareSeqLength, src/io.hpp:352-365; src/subgradient_solver.hpp:112-120)
that all OpenMP threads pull from.  Across GPUs the same order is dealt round-robin: rank r takes entries
r, r+W, r+2W, ... of the cost-sorted list, so every rank gets the same mix of long and short pairs.  Pairs are
independent: there is no exchange step and no collective on the data path; only results are gathered.
"""
import numpy as np


def all_pairs(n):
    ia, ib = np.triu_indices(n, k=1)
    return ia.astype(np.int32), ib.astype(np.int32)


def cost_order(lengths, idx_a, idx_b):
    """Indices of the pairs sorted by DP size (longer sequence, then shorter one), longest first; stable."""
    lengths = np.asarray(lengths, dtype=np.int64)
    la = np.maximum(lengths[idx_a], lengths[idx_b])
    lb = np.minimum(lengths[idx_a], lengths[idx_b])
    return np.argsort(-(la * (1 << 20) + lb), kind="stable")


def shard(order, rank, world):
    """The share of `order` that belongs to `rank` of `world` ranks."""
    if not 0 <= rank < world:
        raise ValueError("rank %d outside world of %d" % (rank, world))
    return order[rank::world]


def shard_cells(lengths, idx_a, idx_b, world):
    """DP cells per rank of the static partition (for load-balance reporting)."""
    lengths = np.asarray(lengths, dtype=np.int64)
    order = cost_order(lengths, idx_a, idx_b)
    cells = lengths[idx_a] * lengths[idx_b]
    return [int(cells[shard(order, r, world)].sum()) for r in range(world)]
