"""jubann_query_sig_noalloc_dev against jubann_query_sig_dev.

The no-alloc entry point exists for hosts that may not call cudaMalloc while serving (a
server that parcels out GPU memory once at start-up, RPC threads with a latency budget --
an allocation stalls the whole device -- or a process where another thread may be recording a
stream).  It runs the same kernels with their scratch in a fixed array inside the library, so
its answers must be the ordinary entry point's answers, bit for bit, call after call even
though the array still holds the previous call's candidates.  Collected last on purpose
(file name): a run that stops at the first failure has then already run every other test.
"""
import numpy as np
import pytest

from gpu_util import random_rows

pytestmark = pytest.mark.gpu


def _run(engine, which, q, qn, k, flags=0):
    import torch
    dev = torch.device("cuda:0")
    nq = q.shape[0]
    d_q = torch.from_numpy(q.view(np.int64)).to(dev)
    d_qn = torch.from_numpy(qn).to(dev) if qn is not None else None
    oi = torch.full((nq, max(k, 1)), -7, dtype=torch.int64, device=dev)
    os_ = torch.full((nq, max(k, 1)), -7.0, dtype=torch.float32, device=dev)
    oc = torch.full((nq,), 12345, dtype=torch.int32, device=dev)
    fn = engine.query_sig_noalloc_dev if which == "noalloc" else engine.query_sig_dev
    fn(nq, d_q.data_ptr(), d_qn.data_ptr() if d_qn is not None else 0, k, flags, oi.data_ptr(), os_.data_ptr(),
       oc.data_ptr(), torch.cuda.current_stream().cuda_stream)
    torch.cuda.synchronize()
    return oi.cpu().numpy(), os_.cpu().numpy(), oc.cpu().numpy()


def _same(a, b):
    for q in range(len(a[2])):
        c = int(b[2][q])
        assert int(a[2][q]) == c
        assert np.array_equal(a[0][q, :c], b[0][q, :c])
        assert np.array_equal(a[1][q, :c].view(np.uint32), b[1][q, :c].view(np.uint32))


@pytest.mark.parametrize("method,hash_num", [("lsh", 512), ("euclid_lsh", 256)])
def test_two_calls_in_a_row_match_the_allocating_entry_point(oracle, method, hash_num):
    from jubatus_b200 import Engine, SIMILAR
    rng = np.random.RandomState(hash_num)
    n, k = 60000, 100
    rows = random_rows(rng, n, hash_num)
    rows[100:400] = rows[7]                         # ties: order must still be by row index
    norms = (rng.rand(n) * 4).astype(np.float32) if method == "euclid_lsh" else None
    with Engine(method, hash_num) as e:
        e.set_rows_sig(np.arange(n), rows, norms)
        # first call: a different, larger batch, so the shared array is full of stale candidates
        q1 = random_rows(rng, 9, hash_num)
        q1[0] = rows[7]
        qn1 = (rng.rand(9) * 4).astype(np.float32) if norms is not None else None
        _same(_run(e, "noalloc", q1, qn1, k), _run(e, "alloc", q1, qn1, k))
        # second call right after it, other queries, fewer of them, other k and flags
        q2 = random_rows(rng, 4, hash_num)
        q2[3] = rows[n - 1]
        qn2 = (rng.rand(4) * 4).astype(np.float32) if norms is not None else None
        _same(_run(e, "noalloc", q2, qn2, 37, SIMILAR), _run(e, "alloc", q2, qn2, 37, SIMILAR))
        # and the first batch again: same answer as before -- and the CPU engine's, checked directly
        again = _run(e, "noalloc", q1, qn1, k)
        _same(again, _run(e, "alloc", q1, qn1, k))
        widx, wscore, wcount, _ = oracle.scan_batch(method, q1, qn1, rows, norms, hash_num, k)
        _same((again[0].view(np.uint64), again[1], again[2]), (widx, wscore, wcount))


def test_small_tables_and_k_edges():
    from jubatus_b200 import Engine
    rng = np.random.RandomState(3)
    hash_num = 64
    q = random_rows(rng, 3, hash_num)
    with Engine("lsh", hash_num) as e:
        for n, k in ((0, 5), (1, 5), (33, 40), (500, 0), (500, 10)):
            e.clear()
            rows = random_rows(rng, n, hash_num)
            if n:
                e.set_rows_sig(np.arange(n), rows)
            _same(_run(e, "noalloc", q, None, k), _run(e, "alloc", q, None, k))


def test_requests_beyond_the_fixed_scratch_are_refused():
    from jubatus_b200 import Engine, JubannError
    rng = np.random.RandomState(5)
    hash_num, n = 64, 300000
    rows = random_rows(rng, n, hash_num)
    with Engine("lsh", hash_num) as e:
        e.set_rows_sig(np.arange(n), rows)
        many = random_rows(rng, 4096, hash_num)     # 4096 queries x chunks x 1000 x 8 B >> 32 MiB
        with pytest.raises(JubannError) as ei:
            _run(e, "noalloc", many, None, 1000)
        assert "scratch" in str(ei.value) and ei.value.code == -1
        with pytest.raises(JubannError) as ei:
            _run(e, "noalloc", many[:2], None, 5000)    # above the single-pass limit
        assert "single-pass" in str(ei.value)
        # the handle is still usable afterwards
        _same(_run(e, "noalloc", many[:8], None, 10), _run(e, "alloc", many[:8], None, 10))
