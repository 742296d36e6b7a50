"""BASELINE.json config 5 through the C ABI entry point orbm_knn2_sharded (include/orbx.h): database sharded by rows,
ncclAllGather of the per-query top-2 records, device merge, chunked over the queries.

  * one GPU: a one-rank NCCL communicator made through orbm_nccl_unique_id / orbm_nccl_comm_create walks the whole
    exchange path (chunks, gather, merge) and must reproduce the plain single-GPU scan;
  * two or more GPUs (skipped otherwise): one process per GPU, each holding one shard; every rank must end with the
    top-2 of the WHOLE database == single-GPU scan == oracle (sampled)."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.multiprocessing as mp

import oracle_lib as O
from orb_slam2_map_b200 import capi, sharding, synth

pytestmark = pytest.mark.gpu


def _single(q, db):
    qd, dd = torch.from_numpy(q).cuda(), torch.from_numpy(db).cuda()
    od = torch.empty((len(q), 2), dtype=torch.int16, device="cuda")
    oi = torch.empty((len(q), 2), dtype=torch.int32, device="cuda")
    capi.knn2_device(qd.data_ptr(), len(q), dd.data_ptr(), len(db), 0, od.data_ptr(), oi.data_ptr(), torch.cuda.current_stream().cuda_stream)
    torch.cuda.synchronize()
    return od.cpu().numpy().view(np.uint16), oi.cpu().numpy()


def test_sharded_entry_point_one_rank_walks_the_exchange():
    nq, ndb = 5000, 30000
    db = synth.random_descriptors(ndb, 8)
    q = synth.plant_near_duplicates(synth.random_descriptors(nq, 7), db, frac=0.02)
    od_w, oi_w = _single(q, db)
    comm = capi.nccl_comm_create(capi.nccl_unique_id(), 0, 1)
    qd, dd = torch.from_numpy(q).cuda(), torch.from_numpy(db).cuda()
    st = torch.cuda.current_stream().cuda_stream
    for chunk in (0, 1024, 2048, 4096, 8192):          # 0: the short cut; the others: 5, 3, 2 and 1 chunks with a ragged tail
        od = torch.zeros((nq, 2), dtype=torch.int16, device="cuda")
        oi = torch.zeros((nq, 2), dtype=torch.int32, device="cuda")
        capi.knn2_sharded(comm, qd.data_ptr(), nq, dd.data_ptr(), ndb, 0, od.data_ptr(), oi.data_ptr(), st, chunk)
        torch.cuda.synchronize()
        assert np.array_equal(oi.cpu().numpy(), oi_w), chunk
        assert np.array_equal(od.cpu().numpy().view(np.uint16), od_w), chunk
    capi.nccl_comm_destroy(comm)
    od_o, oi_o = O.knn2(q[:100], db, nthreads=4)
    assert np.array_equal(oi_w[:100], oi_o) and np.array_equal(od_w[:100], od_o)


def _worker(rank, world, port, nq, ndb, out):
    import torch.distributed as dist
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    torch.cuda.set_device(rank)
    dev = torch.device("cuda", rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev)
    db = synth.random_descriptors(ndb, 18)
    q = synth.plant_near_duplicates(synth.random_descriptors(nq, 17), db, frac=0.02)
    s0, cnt = sharding.shard_rows(ndb, world, rank)
    comm = sharding.ShardComm(dev)
    qd = torch.from_numpy(q).to(dev)
    shard = torch.from_numpy(db[s0:s0 + cnt]).to(dev)
    res = {}
    for chunk in (0, 2048):
        od, oi = sharding.knn2_sharded(qd, shard, s0, comm, query_chunk=chunk)
        torch.cuda.synchronize()
        res[chunk] = (od.cpu().numpy().view(np.uint16), oi.cpu().numpy())
    # the torch.distributed spelling of the same exchange
    od_t, oi_t = sharding.knn2_sharded_torch(qd, shard, s0)
    torch.cuda.synchronize()
    assert np.array_equal(oi_t.cpu().numpy(), res[0][1]) and np.array_equal(od_t.cpu().numpy().view(np.uint16), res[0][0])
    assert np.array_equal(res[0][1], res[2048][1]) and np.array_equal(res[0][0], res[2048][0])
    np.savez(out + f".{rank}.npz", od=res[0][0], oi=res[0][1])
    comm.close()
    dist.barrier()
    dist.destroy_process_group()


def _free_port():
    s = socket.socket(); s.bind(("127.0.0.1", 0)); p = s.getsockname()[1]; s.close()
    return p


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs two GPUs (one process per GPU)")
def test_sharded_entry_point_multi_gpu(tmp_path):
    world = min(torch.cuda.device_count(), 4)
    nq, ndb = 6000, 50001
    out = str(tmp_path / "r")
    mp.spawn(_worker, args=(world, _free_port(), nq, ndb, out), nprocs=world, join=True)
    db = synth.random_descriptors(ndb, 18)
    q = synth.plant_near_duplicates(synth.random_descriptors(nq, 17), db, frac=0.02)
    od_w, oi_w = _single(q, db)
    for r in range(world):
        z = np.load(out + f".{r}.npz")
        assert np.array_equal(z["oi"], oi_w) and np.array_equal(z["od"], od_w), f"rank {r} differs from the single-GPU scan"
