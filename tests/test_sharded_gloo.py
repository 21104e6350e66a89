"""World-size-2 (gloo, CPU) test of the multi-process plumbing that bench.py uses on N GPUs:

* the k-NN database is sharded by rows over the ranks, every rank produces its own top-k lists with GLOBAL row ids, the
  lists are all-gathered and merged with the C ABI's host-side orbm_knn_merge (the (distance, row) tie rule) -- the merged
  result must equal a single-process k-NN over the whole database;
* frames are independent units: rank r owns frames r, r+W, ... and no data-path collective is needed -- only the timing
  reduction (MAX over ranks), which is checked here too.

The per-shard lists come from the CPU oracle (test infrastructure); on the GPU box the same lists come from orbm_knn_device
(tests/test_matcher_gpu.py checks that one against the oracle)."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, ret):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from mam3slam_b200 import ORBmatcher
        from oracle import matcher as om
        rng = np.random.default_rng(7)                        # same data on every rank
        ndb, nq, k = 5000, 64, 3
        db = rng.integers(0, 256, (ndb, 32), dtype=np.uint8)
        db[1234] = db[77]                                      # exact duplicates in different shards: the tie rule decides
        db[4321] = db[77]
        q = db[rng.integers(0, ndb, nq)].copy()
        q[0] = db[77]
        q[:, 5] ^= 3
        lo, hi = ndb * rank // world, ndb * (rank + 1) // world
        d, i = om.knn(q, db[lo:hi], k)
        i = (i + lo).astype(np.int32)                          # global row ids, as orbm_db_create(row_offset=lo) yields
        # distances travel as bytes: neither gloo nor NCCL reduces / gathers 16-bit integers natively (bench.py does the same)
        gd = [torch.empty((nq, 2 * k), dtype=torch.uint8) for _ in range(world)]
        gi = [torch.empty((nq, k), dtype=torch.int32) for _ in range(world)]
        dist.all_gather(gd, torch.from_numpy(d.view(np.uint8)))
        dist.all_gather(gi, torch.from_numpy(i))
        md, mi = ORBmatcher.knn_merge(torch.stack(gd).numpy().view(np.uint16).reshape(world, nq, k), torch.stack(gi).numpy())
        fd, fi = om.knn(q, db, k)
        ok = bool(np.array_equal(md, fd) and np.array_equal(mi, fi))
        # timing reduction used by bench.py: MAX over ranks
        t = torch.tensor([1.0 + rank], dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ok = ok and float(t.item()) == float(world)
        # frame ownership: a partition of the stream, no overlaps, no gaps
        mine = torch.zeros(37, dtype=torch.int32)
        mine[rank::world] = 1
        dist.all_reduce(mine)
        ok = ok and bool((mine == 1).all())
        ret[rank] = ok
    finally:
        dist.destroy_process_group()


@pytest.mark.timeout(180)
def test_sharded_knn_merge_world2():
    world = 2
    port = _free_port()
    with mp.Manager() as man:
        ret = man.dict()
        mp.spawn(_worker, args=(world, port, ret), nprocs=world, join=True)
        assert dict(ret) == {0: True, 1: True}
