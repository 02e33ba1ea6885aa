"""world_size-2 gloo tests (CPU) of the multi-rank host logic: database sharding + all-gather gives every rank all shards'
records in global-row order, merging them gives the single-scan answer, and frame sharding covers every frame exactly
once.  The per-shard searches and the merge are done here by the CPU oracle (test infrastructure); on GPUs the same
plumbing runs over NCCL with k_top2_partial / k_top2_merge (tests/test_gpu_parity.py checks those against the oracle)."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, ndb, nq, strict, ret):
    import sys
    sys.path.insert(0, ROOT)
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from oracle import orb_oracle as O
    from orb_slam3_detailed_comments_kor_b200 import sharding, synth
    db, q = synth.descriptors(77, ndb, nq)
    db[ndb - 5] = db[3]            # duplicate rows living in different shards: the tie must go to the lower global row
    q[0] = db[3]
    lo, hi = sharding.shard_rows(ndb, world, rank)
    part = O.hamming_top2(q, db[lo:hi], 50, 0.9, strict)
    part["idx"] = np.where(part["idx"] >= 0, part["idx"] + lo, -1)
    local = torch.from_numpy(part.view(np.int32).reshape(nq, 4).copy())
    gathered = sharding.gather_top2(local, world)
    parts = gathered.numpy().view(O.MATCH_DTYPE).reshape(world, nq)          # gathered in rank order
    merged = O.merge_top2_host(parts, 50, 0.9, strict).view(np.int32).reshape(nq, 4)   # the merge rule's oracle (on GPUs: k_top2_merge)
    whole = O.hamming_top2(q, db, 50, 0.9, strict).view(np.int32).reshape(nq, 4)
    frames = sharding.frames_for_rank(11, world, rank)
    allf = [None] * world
    dist.all_gather_object(allf, frames)
    ok = bool(np.array_equal(merged, whole)) and merged[0, 2] == 3 and sorted(sum(allf, [])) == list(range(11))
    ret[rank] = ok
    dist.destroy_process_group()


@pytest.mark.parametrize("strict", [False, True])
def test_sharded_top2_merge_over_gloo(strict):
    world = 2
    mgr = mp.Manager()
    ret = mgr.dict()
    port = _free_port()
    mp.spawn(_worker, args=(world, port, 6001, 257, strict, ret), nprocs=world, join=True)
    assert all(ret.get(r) is True for r in range(world)), dict(ret)


def test_shard_ranges_partition_the_database():
    from orb_slam3_detailed_comments_kor_b200 import sharding
    for ndb, world in [(1 << 20, 8), (1000, 3), (7, 2), (5, 8)]:
        spans = [sharding.shard_rows(ndb, world, r) for r in range(world)]
        assert spans[0][0] == 0 and spans[-1][1] == ndb
        assert all(spans[i][1] == spans[i + 1][0] for i in range(world - 1))
