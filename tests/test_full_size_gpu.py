"""BASELINE.json configurations at (or near) their full sizes, checked through size-independent
properties and an independent torch computation of the distances (the oracle would need
minutes per query at these sizes):

* every returned list is ascending in (score, row index) and has min(k, N) entries;
* recomputing the distance of every returned row from the stored signatures gives the score;
* completeness: for a few queries ALL distances are recomputed with plain torch integer ops
  (SWAR popcount on int64) and the exact top-k by (distance, index) must equal the result;
* planted exact duplicates of a query come back first, lowest row index first.
"""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _popcount64(x):
    import torch
    m1, m2, m4 = 0x5555555555555555, 0x3333333333333333, 0x0F0F0F0F0F0F0F0F
    x = x - ((x >> 1) & m1)
    x = (x & m2) + ((x >> 2) & m2)
    x = (x + (x >> 4)) & m4
    # sum the 8 bytes without a multiply that would overflow int64 semantics
    x = x + (x >> 8)
    x = x + (x >> 16)
    x = x + (x >> 32)
    return x & 0x7F


def _torch_hamming(table, q):
    """table [n][w] int64 (device), q [w] int64 -> int32 distances [n] (logical shifts emulated)."""
    import torch
    d = torch.zeros(table.shape[0], dtype=torch.int64, device=table.device)
    for w in range(table.shape[1]):
        x = table[:, w] ^ q[w]
        lo = x & 0xFFFFFFFF
        hi = (x >> 32) & 0xFFFFFFFF
        d += _popcount64(lo) + _popcount64(hi)
    return d


def _fill(engine, n, w64, euclid, seed):
    import torch
    dev = torch.device("cuda:0")
    gen = torch.Generator(device=dev)
    gen.manual_seed(seed)
    stream = torch.cuda.current_stream().cuda_stream
    engine.set_option("reserve_rows", n)
    table = torch.empty((n, w64), dtype=torch.int64, device=dev)
    norms = torch.empty(n, dtype=torch.float32, device=dev) if euclid else None
    chunk = 4_000_000
    for r0 in range(0, n, chunk):
        m = min(chunk, n - r0)
        table[r0:r0 + m] = torch.randint(-2 ** 63, 2 ** 63 - 1, (m, w64), dtype=torch.int64, device=dev, generator=gen)
        if euclid:
            norms[r0:r0 + m] = torch.rand(m, device=dev, generator=gen) * 3 + 0.5
    return table, norms, stream


def _query(engine, q, qn, k, flags=0):
    import torch
    dev = q.device
    nq = q.shape[0]
    oi = torch.zeros((nq, k), dtype=torch.int64, device=dev)
    os_ = torch.zeros((nq, k), dtype=torch.float32, device=dev)
    oc = torch.zeros(nq, dtype=torch.int32, device=dev)
    engine.query_sig_dev(nq, q.data_ptr(), qn.data_ptr() if qn is not None else 0, k, flags, oi.data_ptr(),
                         os_.data_ptr(), oc.data_ptr(), torch.cuda.current_stream().cuda_stream)
    torch.cuda.synchronize()
    return oi, os_, oc


def test_c2_lsh_512_bits_10m_rows_k10():
    """config 2: lsh hash_num=512, 10M rows, query batches, k=10."""
    import torch
    from jubatus_b200 import Engine
    n, h, w64, k, nq = 10_000_000, 512, 8, 10, 256
    with Engine("lsh", h) as e:
        table, _, stream = _fill(e, n, w64, False, 11)
        # plant duplicates: rows 7, 4_000_007 and n-1 are equal
        table[4_000_007] = table[7]
        table[n - 1] = table[7]
        e.set_rows_sig_dev(n, 0, table.data_ptr(), 0, stream)
        gen = torch.Generator(device=table.device); gen.manual_seed(5)
        q = torch.randint(-2 ** 63, 2 ** 63 - 1, (nq, w64), dtype=torch.int64, device=table.device, generator=gen)
        q[0] = table[7]
        q[1] = table[123456]
        oi, os_, oc = _query(e, q, None, k)
        assert int(oc.min()) == k and int(oc.max()) == k
        assert list(oi[0, :3].cpu().numpy()) == [7, 4_000_007, n - 1] and float(os_[0, 2]) == 0.0
        assert int(oi[1, 0]) == 123456 and float(os_[1, 0]) == 0.0
        # ascending (score, index)
        s, i = os_.cpu().numpy(), oi.cpu().numpy()
        assert ((np.diff(s, axis=1) > 0) | ((np.diff(s, axis=1) == 0) & (np.diff(i, axis=1) > 0))).all()
        # completeness on a handful of queries, all distances by torch
        for qi in (0, 1, 2, 100, 255):
            d = _torch_hamming(table, q[qi])
            key = d * n + torch.arange(n, device=d.device)
            top = torch.topk(key, k, largest=False).values
            assert torch.equal(top % n, oi[qi]), qi
            assert np.array_equal((top // n).cpu().numpy().astype(np.float32) / np.float32(h), s[qi])


def test_c3_euclid_256_bits_k100_property_checks():
    """config 3 at its full size on one GPU: euclid_lsh hash_num=256, 10^8 rows, k=100."""
    import torch
    from jubatus_b200 import Engine
    n, h, w64, k, nq = 100_000_000, 256, 4, 100, 32
    with Engine("euclid_lsh", h) as e:
        table, norms, stream = _fill(e, n, w64, True, 12)
        e.set_rows_sig_dev(n, 0, table.data_ptr(), norms.data_ptr(), stream)
        gen = torch.Generator(device=table.device); gen.manual_seed(6)
        q = torch.randint(-2 ** 63, 2 ** 63 - 1, (nq, w64), dtype=torch.int64, device=table.device, generator=gen)
        qn = torch.rand(nq, device=table.device, generator=gen) * 3 + 0.5
        q[0] = table[999]; qn[0] = norms[999]
        oi, os_, oc = _query(e, q, qn, k)
        s = os_.cpu().numpy()
        assert int(oc.min()) == k
        assert (np.diff(s, axis=1) >= 0).all()
        assert int(oi[0, 0]) == 999 and abs(float(os_[0, 0])) < 1e-3
        # recompute every returned distance in float64 from the stored signature and norm
        for qi in (0, 5, 31):
            rows = oi[qi]
            d = _torch_hamming(table[rows], q[qi]).double()
            ni, nq_ = norms[rows].double(), qn[qi].double()
            want = torch.sqrt(torch.clamp(nq_ * nq_ + ni * (ni - 2 * nq_ * torch.cos(d * np.pi / h)), min=0))
            assert torch.allclose(os_[qi].double(), want, rtol=1e-5, atol=2e-3)
            # completeness: nothing outside the list scores below the k-th
            dall = _torch_hamming(table, q[qi]).double()
            nall = norms.double()
            sc = nall * (nall - 2 * nq_ * torch.cos(dall * np.pi / h))
            kth = torch.topk(sc, k, largest=False).values[-1]
            inlist = torch.zeros(n, dtype=torch.bool, device=sc.device); inlist[rows] = True
            assert int(((sc < kth - 1e-6) & ~inlist).sum()) == 0


def test_c4_minhash_128_bits_20m_rows_similar_k50():
    """config 4: minhash hash_num=128, 20M rows, similar_row k=50 (score = 1 - distance)."""
    import torch
    from jubatus_b200 import Engine, SIMILAR
    n, h, w64, k, nq = 20_000_000, 128, 2, 50, 64
    with Engine("minhash", h) as e:
        table, _, stream = _fill(e, n, w64, False, 13)
        e.set_rows_sig_dev(n, 0, table.data_ptr(), 0, stream)
        gen = torch.Generator(device=table.device); gen.manual_seed(7)
        q = torch.randint(-2 ** 63, 2 ** 63 - 1, (nq, w64), dtype=torch.int64, device=table.device, generator=gen)
        q[3] = table[n - 2]
        oi, os_, oc = _query(e, q, None, k, SIMILAR)
        assert int(oc.min()) == k
        assert int(oi[3, 0]) == n - 2 and float(os_[3, 0]) == 1.0
        assert (np.diff(os_.cpu().numpy(), axis=1) <= 0).all()          # similarity descends
        for qi in (3, 10):
            d = _torch_hamming(table, q[qi])
            key = d * n + torch.arange(n, device=d.device)
            top = torch.topk(key, k, largest=False).values
            assert torch.equal(top % n, oi[qi])
            want = np.float32(1.0) - (top // n).cpu().numpy().astype(np.float32) / np.float32(h)
            assert np.array_equal(want, os_[qi].cpu().numpy())


def test_headline_shape_100m_rows_k100_both_regimes():
    """the headline configuration at its full size (lsh 512 bits, 10^8 rows, k=100), in the HBM-bound (2 queries
    per pass) and the popcount-bound (library default) regime: both must give the same exact answer."""
    import torch
    from jubatus_b200 import Engine
    n, h, w64, k, nq = 100_000_000, 512, 8, 100, 32
    with Engine("lsh", h) as e:
        table, _, stream = _fill(e, n, w64, False, 14)
        e.set_rows_sig_dev(n, 0, table.data_ptr(), 0, stream)
        gen = torch.Generator(device=table.device); gen.manual_seed(8)
        q = torch.randint(-2 ** 63, 2 ** 63 - 1, (nq, w64), dtype=torch.int64, device=table.device, generator=gen)
        q[4] = table[n - 1]
        e.set_option("queries_per_pass", 2)
        a = _query(e, q, None, k)
        e.set_option("queries_per_pass", 0)
        b = _query(e, q, None, k)
        assert torch.equal(a[0], b[0]) and torch.equal(a[1], b[1]) and torch.equal(a[2], b[2])
        assert int(a[0][4, 0]) == n - 1
        for qi in (4, 9):
            d = _torch_hamming(table, q[qi])
            key = d * n + torch.arange(n, device=d.device)
            top = torch.topk(key, k, largest=False).values
            assert torch.equal(top % n, a[0][qi])
