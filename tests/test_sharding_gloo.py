"""CPU tests of the multi-process host logic with the gloo backend (world_size 2 and 3): shard plans,
the all_gather layout of the 2-NN exchange, and sharded == whole (oracle per shard + reference merge)."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

import oracle_lib as O
from orb_slam2_map_b200 import sharding, synth


def test_shard_plans_cover_everything():
    for n in (0, 1, 7, 4096, 16_000_001):
        for world in (1, 2, 3, 8):
            got = []
            for r in range(world):
                s, c = sharding.shard_rows(n, world, r)
                got.append((s, c))
            assert got[0][0] == 0 and sum(c for _, c in got) == n
            for (s0, c0), (s1, _) in zip(got, got[1:]):
                assert s0 + c0 == s1
    # consecutive-frame matching: every pair (k, k+1) is owned by exactly one rank
    for n in (2, 9, 4096):
        for world in (1, 2, 4, 8):
            pairs = []
            for r in range(world):
                s, nf, npairs = sharding.frame_shard_with_halo(n, world, r)
                assert npairs == max(nf - 1, 0) or nf == 0
                pairs += [(s + i, s + i + 1) for i in range(npairs)]
            assert pairs == [(k, k + 1) for k in range(n - 1)]


def _merge_numpy(pd, pi):
    """reference merge: two smallest (distance, global index) per query over all parts"""
    world, nq, _ = pd.shape
    od = np.empty((nq, 2), np.uint16); oi = np.empty((nq, 2), np.int32)
    for q in range(nq):
        c = sorted((int(pd[p, q, j]), int(pi[p, q, j])) for p in range(world) for j in range(2) if pi[p, q, j] >= 0)
        c += [(257, -1)] * 2
        od[q] = (c[0][0], c[1][0]); oi[q] = (c[0][1], c[1][1])
    return od, oi


def _worker(rank, world, port, nq, ndb, out):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    q = synth.plant_near_duplicates(synth.random_descriptors(nq, 5), synth.random_descriptors(ndb, 6), frac=0.05)
    db = synth.random_descriptors(ndb, 6)
    start, count = sharding.shard_rows(ndb, world, rank)
    ld, li = O.knn2(q, db[start:start + count], offset=start)          # the per-rank compute (oracle on CPU)
    pd, pi = sharding.gather_top2(torch.from_numpy(ld.view(np.int16)), torch.from_numpy(li))
    assert pd.shape == (world, nq, 2) and pi.shape == (world, nq, 2)
    # part r must be rank r's local result on every rank
    s2, c2 = sharding.shard_rows(ndb, world, (rank + 1) % world)
    od2, oi2 = O.knn2(q, db[s2:s2 + c2], offset=s2)
    assert np.array_equal(pi[(rank + 1) % world].numpy(), oi2)
    od, oi = _merge_numpy(pd.numpy().view(np.uint16), pi.numpy())
    wd, wi = O.knn2(q, db)
    assert np.array_equal(oi, wi) and np.array_equal(od, wd)
    if rank == 0:
        np.save(out, oi)
    dist.barrier()
    dist.destroy_process_group()


def _free_port():
    s = socket.socket(); s.bind(("127.0.0.1", 0)); p = s.getsockname()[1]; s.close()
    return p


@pytest.mark.parametrize("world", [2, 3])
def test_knn2_sharded_exchange_gloo(world, tmp_path):
    out = str(tmp_path / "oi.npy")
    mp.spawn(_worker, args=(world, _free_port(), 257, 5001, out), nprocs=world, join=True)
    assert np.load(out).shape == (257, 2)
