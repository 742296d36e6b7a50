"""Multi-GPU partitioning of the hot path.  One process per GPU.

* frames / stereo pairs are independent units: contiguous ranges per rank, no data-path collective;
  consecutive-frame matching (SearchForInitialization on pairs (k, k+1)) needs a ONE-frame halo so
  that the pair straddling a shard boundary is local to the lower rank.
* brute-force 2-NN (config 5): the database is sharded by rows, queries are replicated; every rank
  produces its local (best, second) per query with GLOBAL row indices and the ranks exchange them
  with one all_gather (12 bytes per query and rank); the merge keeps the two smallest
  (distance, global index) pairs — the same order the single-GPU scan produces.

Host logic only (torch.distributed); works with the nccl backend on GPUs and gloo on CPU (tests).
"""
from __future__ import annotations

from typing import Tuple


def shard_rows(n: int, world: int, rank: int) -> Tuple[int, int]:
    """Contiguous split of n units: (start, count) of `rank`; the first n % world ranks get one more."""
    base, rem = divmod(n, world)
    start = rank * base + min(rank, rem)
    return start, base + (1 if rank < rem else 0)


def frame_shard_with_halo(nframes: int, world: int, rank: int) -> Tuple[int, int, int]:
    """(first frame, frames to extract, pairs to match) for consecutive-frame matching.
    Rank r owns pairs (k, k+1) for k in its frame range, so it also extracts the first frame of the
    next shard (duplicated work: one frame per boundary)."""
    start, count = shard_rows(nframes, world, rank)
    if count == 0:
        return start, 0, 0
    last = start + count            # exclusive
    halo = 1 if last < nframes else 0
    return start, count + halo, count - 1 + halo


def gather_top2(dist, idx, group=None):
    """all_gather of per-rank local top-2 results.  dist: [nq,2] int16 (uint16 bit patterns),
    idx: [nq,2] int32 -> (parts_dist [world,nq,2], parts_idx [world,nq,2]) on every rank."""
    import torch
    import torch.distributed as dist_
    world = dist_.get_world_size(group)
    nq = dist.shape[0]
    # one exchange of 12 bytes per query: (d1, d2 : u16) + (i1, i2 : i32), as bytes (every backend moves uint8)
    rec = torch.cat([dist.contiguous().view(torch.uint8).reshape(nq, 4), idx.contiguous().view(torch.uint8).reshape(nq, 8)], dim=1)
    allrec = torch.empty((world * nq, 12), dtype=torch.uint8, device=rec.device)
    dist_.all_gather_into_tensor(allrec, rec, group=group)               # rank-major concatenation
    allrec = allrec.view(world, nq, 12)
    pd = allrec[:, :, :4].contiguous().view(dist.dtype).reshape(world, nq, 2)
    pi = allrec[:, :, 4:].contiguous().view(idx.dtype).reshape(world, nq, 2)
    return pd, pi


class ShardComm:
    """The NCCL communicator orbm_knn2_sharded runs on, made through the C ABI (orbm_nccl_unique_id /
    orbm_nccl_comm_create): rank 0 draws the id, torch.distributed only carries its 128 bytes to the other ranks."""

    def __init__(self, device, group=None):
        import torch
        import torch.distributed as dist_
        from . import capi
        self.rank, self.world = dist_.get_rank(group), dist_.get_world_size(group)
        uid = torch.zeros(128, dtype=torch.uint8, device=device)
        if self.rank == 0:
            uid.copy_(torch.frombuffer(bytearray(capi.nccl_unique_id()), dtype=torch.uint8))
        dist_.broadcast(uid, src=0, group=group)
        with torch.cuda.device(device):
            self.comm = capi.nccl_comm_create(bytes(uid.cpu().numpy().tobytes()), self.rank, self.world)

    def close(self):
        from . import capi
        if self.comm:
            capi.nccl_comm_destroy(self.comm)
            self.comm = None


def knn2_sharded(queries, db_shard, shard_offset: int, comm: "ShardComm | None" = None, stream=None, query_chunk=0, out=None):
    """Config 5 on one rank through ONE C-ABI call (orbm_knn2_sharded): local brute-force top-2 on the CUDA shard,
    ncclAllGather of the 12-byte records and the device merge, chunked so that the exchange of a chunk overlaps the scan of
    the next.  queries [nq,32] uint8 cuda, db_shard [n,32] uint8 cuda -> (dist [nq,2] int16, idx [nq,2] int32)."""
    import torch
    from . import capi
    nq = queries.shape[0]
    st = stream if stream is not None else torch.cuda.current_stream().cuda_stream
    od, oi = out if out is not None else (torch.empty((nq, 2), dtype=torch.int16, device=queries.device),
                                          torch.empty((nq, 2), dtype=torch.int32, device=queries.device))
    capi.knn2_sharded(comm.comm if comm is not None else None, queries.data_ptr(), nq, db_shard.data_ptr(), db_shard.shape[0],
                      shard_offset, od.data_ptr(), oi.data_ptr(), st, query_chunk)
    return od, oi


def knn2_sharded_torch(queries, db_shard, shard_offset: int, group=None, stream=None):
    """The same exchange spelled with torch.distributed (any backend): used by the CPU (gloo) tests of the merge order and
    as the cross-check of orbm_knn2_sharded on GPUs."""
    import torch
    from . import capi
    nq = queries.shape[0]
    st = stream if stream is not None else torch.cuda.current_stream().cuda_stream
    ld = torch.empty((nq, 2), dtype=torch.int16, device=queries.device)
    li = torch.empty((nq, 2), dtype=torch.int32, device=queries.device)
    capi.knn2_device(queries.data_ptr(), nq, db_shard.data_ptr(), db_shard.shape[0], shard_offset, ld.data_ptr(),
                     li.data_ptr(), st)
    pd, pi = gather_top2(ld, li, group)
    od = torch.empty_like(ld)
    oi = torch.empty_like(li)
    capi.knn2_merge_device(pd.data_ptr(), pi.data_ptr(), pd.shape[0], nq, od.data_ptr(), oi.data_ptr(), st)
    return od, oi
