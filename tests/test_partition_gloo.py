"""CPU tests of the multi-GPU host logic: the static pair partition, and a world_size-2 `gloo` run in which each
rank solves its shard (with the CPU restatement standing in for the device) and rank 0 gathers and checks that
the union equals the single-rank result."""
import os
import sys

import numpy as np
import pytest

from tests import util

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_shards_are_a_balanced_partition():
    from lara_b200 import partition
    rng = np.random.default_rng(5)
    lengths = np.clip(np.rint(np.exp(rng.normal(np.log(150), 0.6, size=200))), 50, 600).astype(int)
    ia, ib = partition.all_pairs(len(lengths))
    order = partition.cost_order(lengths, ia, ib)
    la = np.maximum(lengths[ia], lengths[ib])[order]
    assert np.all(np.diff(la) <= 0)                                    # longest first
    for world in (1, 2, 4, 8):
        shards = [partition.shard(order, r, world) for r in range(world)]
        allidx = np.concatenate(shards)
        assert len(allidx) == len(ia) and len(np.unique(allidx)) == len(ia)
        cells = partition.shard_cells(lengths, ia, ib, world)
        assert max(cells) / (sum(cells) / world) < 1.01                 # within 1 % of perfect balance
    with pytest.raises(ValueError):
        partition.shard(order, 2, 2)


def _worker(rank, world, port, path, out):
    import torch.distributed as dist
    sys.path.insert(0, ROOT)
    import lara_b200 as lb
    from lara_b200 import partition
    from oracle import binding as ob
    dist.init_process_group("gloo", init_method="tcp://127.0.0.1:%d" % port, rank=rank, world_size=world)
    store = lb.InputStorage.from_files(in_file=path)
    p = lb.Parameters()
    p.set(num_iterations=30)
    ia, ib = partition.all_pairs(len(store))
    order = partition.cost_order(store.lengths(), ia, ib)
    mine = partition.shard(order, rank, world)
    res, _ = ob.solve_pairs(p, store.flat(), ia[mine], ib[mine])
    keys = [(int(mine[k]),) + r.key() for k, r in enumerate(lb.results_to_list(res))]
    ob.free_results(res)
    gathered = [None] * world
    dist.all_gather_object(gathered, keys)
    if rank == 0:
        merged = sorted(sum(gathered, []))
        full, _ = ob.solve_pairs(p, store.flat(), ia, ib)
        want = [(k,) + r.key() for k, r in enumerate(lb.results_to_list(full))]
        ob.free_results(full)
        with open(out, "w") as f:
            f.write("ok" if merged == want else "mismatch")
    dist.barrier()
    dist.destroy_process_group()


def test_two_rank_gloo_gather_equals_single_rank(built, tmp_path):
    import torch.multiprocessing as mp
    files = util.make_input(seed=41, n=6, length=40, lenmax=70)
    out = str(tmp_path / "result.txt")
    port = 29500 + (os.getpid() % 2000)
    mp.spawn(_worker, args=(2, port, files[0], out), nprocs=2, join=True)
    assert open(out).read() == "ok"
