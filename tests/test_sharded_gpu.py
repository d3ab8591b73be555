"""Row sharding + candidate merge: G shard handles on ONE GPU stand in for G ranks (the
all-gather is a torch.cat here; tests/test_dist_gloo.py covers the collective plumbing).
The merged result must be bit-identical to the single-table ranking, ties included."""
import numpy as np
import pytest

from gpu_util import assert_same_ranking, random_rows

pytestmark = pytest.mark.gpu


def _sharded_query(method, hash_num, G, rows, norms, q, qn, k, flags=0):
    import torch
    from jubatus_b200 import Engine
    dev = torch.device("cuda:0")
    n, nq = rows.shape[0], q.shape[0]
    engines = [Engine(method, hash_num, 0, r, G) for r in range(G)]
    try:
        for e in engines:
            if n:
                e.set_rows_sig(np.arange(n), rows, norms)    # every shard sees every row, keeps its own
            assert e.size() == (n, len(range(e.shard_rank, n, G)))
        d_q = torch.from_numpy(q.view(np.int64)).to(dev)
        d_qn = torch.from_numpy(qn).to(dev) if qn is not None else None
        keys, idxs = [], []
        stream = torch.cuda.current_stream().cuda_stream
        for e in engines:
            ck = torch.empty((nq, k), dtype=torch.int32, device=dev)
            ci = torch.empty((nq, k), dtype=torch.int64, device=dev)
            e.scan_local_dev(nq, d_q.data_ptr(), d_qn.data_ptr() if d_qn is not None else 0, k, ck.data_ptr(),
                             ci.data_ptr(), stream)
            keys.append(ck)
            idxs.append(ci)
        all_k = torch.stack(keys).contiguous()      # [G][nq][k] -- what all_gather_into_tensor yields
        all_i = torch.stack(idxs).contiguous()
        out_i = torch.zeros((nq, k), dtype=torch.int64, device=dev)
        out_s = torch.zeros((nq, k), dtype=torch.float32, device=dev)
        out_c = torch.zeros(nq, dtype=torch.int32, device=dev)
        engines[0].merge_dev(nq, k, G, all_k.data_ptr(), all_i.data_ptr(), d_qn.data_ptr() if d_qn is not None else 0,
                             flags, out_i.data_ptr(), out_s.data_ptr(), out_c.data_ptr(), stream)
        torch.cuda.synchronize()
        return (out_i.cpu().numpy().view(np.uint64), out_s.cpu().numpy(), out_c.cpu().numpy().view(np.uint32))
    finally:
        for e in engines:
            e.close()


@pytest.mark.parametrize("G", [2, 3, 8])
def test_sharded_hamming_equals_single_table(oracle, G):
    rng = np.random.RandomState(G)
    hash_num, n, nq, k = 128, 10007, 6, 100
    base = random_rows(rng, 50, hash_num)
    rows = base[rng.randint(0, 50, size=n)]           # heavy ties across shards
    q = np.concatenate([base[:3], random_rows(rng, nq - 3, hash_num)])
    got = _sharded_query("lsh", hash_num, G, rows, None, q, None, k)
    want = oracle.scan_batch("lsh", q, None, rows, None, hash_num, k)[:3]
    assert_same_ranking(got, want, nq, k)


def test_sharded_small_tables(oracle):
    rng = np.random.RandomState(1)
    hash_num, nq, k = 64, 3, 10
    q = random_rows(rng, nq, hash_num)
    for n in (0, 1, 3, 9, 25):
        rows = random_rows(rng, n, hash_num)
        got = _sharded_query("lsh", hash_num, 4, rows, None, q, None, k)
        want = oracle.scan_batch("lsh", q, None, rows, None, hash_num, k)[:3]
        assert_same_ranking(got, want, nq, k)


def test_sharded_euclid(oracle):
    from jubatus_b200 import SIMILAR
    rng = np.random.RandomState(2)
    hash_num, n, nq, k = 256, 8000, 5, 100
    rows = random_rows(rng, n, hash_num)
    norms = (rng.rand(n) * 5).astype(np.float32)
    q = random_rows(rng, nq, hash_num)
    qn = (rng.rand(nq) * 5).astype(np.float32)
    got = _sharded_query("euclid_lsh", hash_num, 4, rows, norms, q, qn, k, SIMILAR)
    want = oracle.scan_batch("euclid_lsh", q, qn, rows, norms, hash_num, k, similar=True)[:3]
    assert_same_ranking(got, want, nq, k)


def test_diff_pack_apply_roundtrip(oracle):
    # MIX transport: rows packed on one replica, applied on another
    import torch
    from jubatus_b200 import Engine
    rng = np.random.RandomState(3)
    hash_num, n = 256, 500
    rows = random_rows(rng, n, hash_num)
    norms = rng.rand(n).astype(np.float32)
    dev = torch.device("cuda:0")
    with Engine("euclid_lsh", hash_num) as a, Engine("euclid_lsh", hash_num) as b:
        a.set_rows_sig(np.arange(n), rows, norms)
        b.set_rows_sig(np.arange(n), np.zeros_like(rows), np.zeros_like(norms))
        changed = np.array([3, 77, 499, 0], dtype=np.int64)
        d_idx = torch.from_numpy(changed).to(dev)
        rb = a.diff_record_bytes()
        assert rb == 8 + 8 * 4 + 8
        rec = torch.zeros(len(changed) * rb, dtype=torch.uint8, device=dev)
        s = torch.cuda.current_stream().cuda_stream
        a.diff_pack_dev(len(changed), d_idx.data_ptr(), rec.data_ptr(), s)
        b.diff_apply_dev(len(changed), rec.data_ptr(), s)
        torch.cuda.synchronize()
        got, gn = b.get_rows_sig(changed)
        other, on = b.get_rows_sig([1, 2])
    assert np.array_equal(got, rows[changed]) and np.array_equal(gn, norms[changed])
    assert (other == 0).all() and (on == 0).all()
