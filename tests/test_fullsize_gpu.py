"""BASELINE.json's full sizes through size-independent properties (the oracle only checks a sample):
configs[1] 4096 x 640x480 frames + consecutive-frame SearchForInitialization; configs[4] per-GPU shard of the
1M x 16M brute-force 2-NN (2M database rows per GPU); configs[3] 1920x1080 / 5000 features."""
import zlib

import numpy as np
import pytest
import torch

import oracle_lib as O
from orb_slam2_map_b200 import capi, synth

pytestmark = pytest.mark.gpu


def _crc(*arrays):
    c = 0
    for a in arrays:
        c = zlib.crc32(np.ascontiguousarray(a).tobytes(), c)
    return c


def test_config1_4096_frames_periodic_batch_and_sampled_oracle():
    """A 4096-frame batch built from 64 distinct frames: every slot must equal the slot 64 earlier (a checksum of
    checksums over the whole batch: batch position, chunk boundaries and ramped chunk sizes must not matter), pair
    matches likewise, and a sample of slots must equal the oracle."""
    F, P, W, H = 4096, 64, 640, 480
    seq = synth.frame_sequence(P, W, H, seed=11)
    host = torch.empty((F, H, W), dtype=torch.uint8, pin_memory=True)
    host.copy_(torch.from_numpy(np.ascontiguousarray(np.tile(seq, (F // P, 1, 1)))))
    ex = capi.Extractor(max_width=W, max_height=H, chunk_frames=512, result_slots=F)
    cap = ex.cap
    kp = np.zeros((F, cap), capi.KP_DTYPE); de = np.zeros((F, cap, 32), np.uint8); cnt = np.zeros(F, np.int32)
    m12 = np.zeros((F - 1, cap), np.int32); nm = np.zeros(F - 1, np.int32)
    ex.extract_match_to_host(host.numpy(), kp, de, cnt, m12, nm, 100, 0.9, True)
    assert cnt.min() > 800 and cnt.max() <= cap
    sums = np.array([_crc(kp[f, :cnt[f]], de[f, :cnt[f]]) for f in range(F)], np.uint32).reshape(F // P, P)
    assert (sums == sums[0]).all()
    # pairs inside a period (the pair P-1 -> P wraps to frame 0 of the next period and is periodic too)
    msum = np.array([_crc(m12[f, :cnt[f]]) for f in range(F - 1)], np.uint32)
    assert (msum[P:] == msum[:-P]).all() and (nm[P:] == nm[:-P]).all()
    orc = O.Extractor()
    for f in (0, 17, 63):
        ko, do = orc.extract(seq[f])
        assert len(ko) == cnt[f] and ko.tobytes() == kp[f, :cnt[f]].tobytes() and np.array_equal(do, de[f, :cnt[f]])
    fa = O.FrameData(kp[5, :cnt[5]], de[5, :cnt[5]], W, H); fb = O.FrameData(kp[6, :cnt[6]], de[6, :cnt[6]], W, H)
    n_o, m_o, _, _ = O.search_for_initialization(fa, fb, np.stack([kp[5, :cnt[5]]["x"], kp[5, :cnt[5]]["y"]], 1), 100, 0.9, True)
    assert n_o == nm[5] and np.array_equal(m_o, m12[5, :cnt[5]])


def test_config4_knn_shard_1m_x_2m_properties():
    """One GPU's share of configs[4]: 1M queries against 2M database rows.  Properties: planted exact copies come
    back with distance 0 at the planted row; d1 <= d2; first index < second on ties; a sample equals the oracle;
    searching the two halves and merging equals the whole (linearity of the top-2 merge)."""
    nq, ndb = 1 << 20, 2_000_000
    g = torch.Generator(device="cuda").manual_seed(7)
    db = torch.randint(0, 256, (ndb, 32), dtype=torch.uint8, device="cuda", generator=g)
    q = torch.randint(0, 256, (nq, 32), dtype=torch.uint8, device="cuda", generator=g)
    planted_q = torch.arange(0, nq, 997, device="cuda")
    planted_row = (planted_q * 7919) % ndb
    q[planted_q] = db[planted_row]
    od = torch.empty((nq, 2), dtype=torch.int16, device="cuda"); oi = torch.empty((nq, 2), dtype=torch.int32, device="cuda")
    st = torch.cuda.current_stream().cuda_stream
    capi.knn2_device(q.data_ptr(), nq, db.data_ptr(), ndb, 0, od.data_ptr(), oi.data_ptr(), st)
    torch.cuda.synchronize()
    d = od.to(torch.int32) & 0xffff
    assert (d[planted_q, 0] == 0).all() and (oi[planted_q, 0].long() == planted_row).all()
    assert (d[:, 0] <= d[:, 1]).all()
    tie = d[:, 0] == d[:, 1]
    assert (oi[tie, 0] < oi[tie, 1]).all()
    assert (oi >= 0).all() and (oi < ndb).all() and (oi[:, 0] != oi[:, 1]).all()
    sample = torch.arange(0, nq, nq // 24, device="cuda")[:24]
    dbh = db.cpu().numpy()
    od_o, oi_o = O.knn2(q[sample].cpu().numpy(), dbh, offset=0, nthreads=16)
    assert np.array_equal(oi[sample].cpu().numpy(), oi_o) and np.array_equal(d[sample].cpu().numpy().astype(np.uint16), od_o.astype(np.uint16))
    # halves + merge == whole
    half = ndb // 2
    pd = torch.empty((2, nq, 2), dtype=torch.int16, device="cuda"); pi = torch.empty((2, nq, 2), dtype=torch.int32, device="cuda")
    for s in range(2):
        capi.knn2_device(q.data_ptr(), nq, db[s * half:].data_ptr(), half, s * half, pd[s].data_ptr(), pi[s].data_ptr(), st)
    md = torch.empty_like(od); mi = torch.empty_like(oi)
    capi.knn2_merge_device(pd.data_ptr(), pi.data_ptr(), 2, nq, md.data_ptr(), mi.data_ptr(), st)
    torch.cuda.synchronize()
    assert torch.equal(md, od) and torch.equal(mi, oi)


def test_config3_hd_5000_features_batch_vs_single_and_oracle():
    W, H, n = 1920, 1080, 6
    seq = synth.frame_sequence(n, W, H, seed=4)
    ex = capi.Extractor(nfeatures=5000, max_width=W, max_height=H, chunk_frames=4, result_slots=n)
    ex.extract_batch(seq)
    k, d, c = ex.download(0, n)
    assert c.min() > 4000
    ko, do = O.Extractor(nfeatures=5000).extract(seq[3])
    assert len(ko) == c[3] and ko.tobytes() == k[3, :c[3]].tobytes() and np.array_equal(do, d[3, :c[3]])
    ex1 = capi.Extractor(nfeatures=5000, max_width=W, max_height=H, chunk_frames=1, result_slots=1)
    for f in (0, 5):
        k1, d1 = ex1.extract(seq[f])
        assert len(k1) == c[f] and k1.tobytes() == k[f, :c[f]].tobytes() and np.array_equal(d1, d[f, :c[f]])
