"""orbm_knn2_device_noalloc (no counterpart in the reference): the brute-force
top-2 search with its per-split scratch in a built-in 8 MiB device array instead of a cudaMallocAsync block.  Checked through
the C ABI against the existing orbm_knn2_device, bit for bit (integer work: tolerance 0), on two calls in a row that share the
array -- the second with other sizes, so stale contents of the first would show.  Collected last."""
import numpy as np
import pytest
import torch

from orb_slam2_map_b200 import capi, synth

pytestmark = pytest.mark.gpu


def _run(fn, q, db, offset):
    qd, dd = torch.from_numpy(q).cuda(), torch.from_numpy(db).cuda()
    od = torch.full((len(q), 2), -1, dtype=torch.int16, device="cuda")
    oi = torch.full((len(q), 2), -7, dtype=torch.int32, device="cuda")
    fn(qd.data_ptr(), len(q), dd.data_ptr(), len(db), offset, od.data_ptr(), oi.data_ptr(), torch.cuda.current_stream().cuda_stream)
    torch.cuda.synchronize()
    return od.cpu().numpy().view(np.uint16), oi.cpu().numpy()


def test_knn2_noalloc_equals_knn2_two_calls_in_a_row():
    capi.lib().orbm_knn2_set_variant(1)
    # few query tiles against many rows: both entry points split the database, so the scratch array is in use
    for nq, ndb, offset, seed in [(3000, 300000, 1000, 1), (1111, 70000, 0, 2)]:
        q = synth.random_descriptors(nq, seed)
        db = synth.random_descriptors(ndb, seed + 10)
        db[700] = db[20]
        q = synth.plant_near_duplicates(q, db, frac=0.05)
        q[5] = db[20]                                            # a tie between two rows: the lower index wins
        od_a, oi_a = _run(capi.knn2_device, q, db, offset)
        od_b, oi_b = _run(capi.knn2_device_noalloc, q, db, offset)
        assert np.array_equal(oi_b, oi_a) and np.array_equal(od_b, od_a)
        assert oi_a[5, 0] == offset + 20 and od_a[5, 0] == 0


def test_knn2_noalloc_rejects_sizes_beyond_the_array():
    # 9 splits of 2^23 rows are needed: more than the array's 8.  Nothing is launched, so the pointers are never read.
    buf = torch.zeros(64, dtype=torch.uint8, device="cuda")
    with pytest.raises(capi.OrbxError) as e:
        capi.knn2_device_noalloc(buf.data_ptr(), 16, buf.data_ptr(), 8 * (1 << 23) + 1, 0, buf.data_ptr(), buf.data_ptr())
    assert "scratch" in str(e.value)
