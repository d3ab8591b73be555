"""Table semantics at the C-ABI that the reference's column_table gives for free and a batched
device table has to work for: a batch that names a row twice, a failed batch, bits above hash_num,
and the order of device-pointer writes against the handle's own stream."""
import numpy as np
import pytest

from gpu_util import assert_same_ranking, random_rows

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("method,hash_num", [("lsh", 512), ("euclid_lsh", 256), ("lsh", 64)])
def test_a_batch_that_names_a_row_twice_stores_the_last_entry(method, hash_num):
    # column_table::add applied in order: the later entry of the batch is what stays
    from jubatus_b200 import Engine
    rng = np.random.RandomState(1)
    n = 3000
    idx = rng.randint(0, 40, size=n).astype(np.uint64)         # 3000 entries onto 40 rows
    idx[:40] = np.arange(40)                                    # (each row appended before it is overwritten)
    rows = random_rows(rng, n, hash_num)
    norms = rng.rand(n).astype(np.float32) if method == "euclid_lsh" else None
    last = {int(g): j for j, g in enumerate(idx)}
    with Engine(method, hash_num) as e:
        e.set_rows_sig(idx, rows, norms)
        assert e.size()[0] == 40
        back, bn = e.get_rows_sig(np.arange(40, dtype=np.uint64))
    for g in range(40):
        assert np.array_equal(back[g], rows[last[g]]), g        # whole rows, never a mix of two entries
        if norms is not None:
            assert bn[g] == norms[last[g]]


def test_a_csr_batch_that_names_a_row_twice(feature_tables):
    from jubatus_b200 import Engine, workload as w
    s32, k64 = feature_tables
    ids, vals = w.synth_rows(w.MASTER_SEED, 0, 6)
    csr = w.to_csr(ids, vals, s32, k64)
    with Engine("lsh", 128) as e:
        sigs, _ = e.hash_csr(*csr)
        e.set_rows_csr(np.array([0, 1, 0, 2, 1, 0], dtype=np.uint64), *csr)
        assert e.size()[0] == 3
        back, _ = e.get_rows_sig(np.arange(3, dtype=np.uint64))
    assert np.array_equal(back, sigs[[5, 4, 3]])


def test_a_refused_batch_leaves_the_table_as_it_was():
    from jubatus_b200 import Engine, JubannError
    rng = np.random.RandomState(2)
    rows = random_rows(rng, 10, 128)
    with Engine("lsh", 128) as e:
        e.set_rows_sig(np.arange(4, dtype=np.uint64), rows[:4])
        with pytest.raises(JubannError) as ei:
            e.set_rows_sig(np.array([4, 5, 9], dtype=np.uint64), rows[4:7])     # 9 skips past the end
        assert ei.value.code == -5
        assert e.size() == (4, 4)
        back, _ = e.get_rows_sig(np.arange(4, dtype=np.uint64))
        assert np.array_equal(back, rows[:4])
        got = e.query_sig(rows[:1], None, 10)
        assert got[2][0] == 4


@pytest.mark.parametrize("method,hash_num", [("lsh", 100), ("euclid_lsh", 100), ("minhash", 33), ("lsh", 1000)])
def test_bits_at_or_above_hash_num_are_ignored(oracle, method, hash_num):
    # storage::bit_vector keeps the unused bits of its last word zero; words handed in with garbage
    # there must rank exactly like the clean ones (a stray bit would push a distance past hash_num)
    from jubatus_b200 import Engine
    rng = np.random.RandomState(5)
    n, nq, k = 4000, 5, 20
    rows = random_rows(rng, n, hash_num)
    q = np.concatenate([rows[:2], random_rows(rng, nq - 2, hash_num)])
    norms = (rng.rand(n) * 3).astype(np.float32) if method == "euclid_lsh" else None
    qn = (rng.rand(nq) * 3).astype(np.float32) if method == "euclid_lsh" else None
    garbage = ~np.uint64((1 << (hash_num % 64)) - 1)
    dirty_rows, dirty_q = rows.copy(), q.copy()
    dirty_rows[:, -1] |= garbage
    dirty_q[:, -1] |= garbage
    with Engine(method, hash_num) as e:
        e.set_rows_sig(np.arange(n), dirty_rows, norms)
        back, _ = e.get_rows_sig(np.arange(n))
        assert np.array_equal(back, rows)                       # stored clean
        got = e.query_sig(dirty_q, qn, k)
    want = oracle.scan_batch(method, q, qn, rows, norms, hash_num, k)[:3]
    assert_same_ranking(got, want, nq, k)


def test_device_pointer_writes_are_ordered_before_the_handle_stream(oracle):
    # rows written through a _dev entry point on a side stream, no reserve, no synchronisation: the
    # next host-pointer call (which grows the slab and scans on the handle's own stream) sees them
    import torch
    from jubatus_b200 import Engine
    rng = np.random.RandomState(7)
    hash_num, k = 256, 10
    dev = torch.device("cuda:0")
    side = torch.cuda.Stream(dev)
    total = 0
    chunks = [random_rows(rng, n, hash_num) for n in (3000, 50000, 200000, 7)]
    q = random_rows(rng, 3, hash_num)
    with Engine("lsh", hash_num) as e:
        keep = []
        for rows in chunks:
            with torch.cuda.stream(side):
                d = torch.from_numpy(rows.view(np.int64)).to(dev, non_blocking=True)
                keep.append(d)
                e.set_rows_sig_dev(len(rows), total, d.data_ptr(), 0, side.cuda_stream)
            total += len(rows)
            allrows = np.concatenate(chunks[: len(keep)])
            got = e.query_sig(np.concatenate([q, allrows[-1:]]), None, k)     # right away, on the handle's stream
            want = oracle.scan_batch("lsh", np.concatenate([q, allrows[-1:]]), None, allrows, None, hash_num, k)[:3]
            assert_same_ranking(got, want, 4, k)
        back, _ = e.get_rows_sig(np.arange(total, dtype=np.uint64))
        assert np.array_equal(back, np.concatenate(chunks))


@pytest.mark.parametrize("method,hash_num", [("lsh", 256), ("euclid_lsh", 256), ("minhash", 128)])
def test_set_row_query_is_set_row_then_neighbor_row_from_id(oracle, feature_tables, method, hash_num):
    # jubann_set_row_query_csr: one call stores a row and ranks the table with the row in it -- the answers of
    # set_rows_csr followed by query_rows, for appended rows and for rows overwritten in place
    from jubatus_b200 import Engine, SIMILAR, workload as w
    s32, k64 = feature_tables
    n, k = 300, 12
    ids, vals = w.synth_rows(w.MASTER_SEED, 0, n + 20, nnz=30, nfeat_log2=10)
    euclid = method == "euclid_lsh"
    with Engine(method, hash_num) as fused, Engine(method, hash_num) as plain:
        csr = w.to_csr(ids[:n], vals[:n], s32, k64)
        fused.set_rows_csr(np.arange(n), *csr)
        plain.set_rows_csr(np.arange(n), *csr)
        for step, (src, at) in enumerate([(n, n), (n + 1, n + 1), (n + 2, 5), (n + 3, n + 2), (n + 4, 0)]):
            indptr, seed32, keyhash, v = w.to_csr(ids[src:src + 1], vals[src:src + 1], s32, k64)
            flags = SIMILAR if step % 2 else 0
            got = fused.set_row_query_csr(at, seed32, keyhash, v, k, flags)
            plain.set_rows_csr(np.array([at], dtype=np.uint64), indptr, seed32, keyhash, v)
            want = plain.query_rows(np.array([at], dtype=np.uint64), k, flags)
            assert_same_ranking(got, want, 1, k)
            assert got[0][0][0] == at or got[1][0][0] == want[1][0][0]      # the row itself leads (ties by row index aside)
            assert fused.size() == plain.size()
        total = fused.size()[0]
        a, an = fused.get_rows_sig(np.arange(total, dtype=np.uint64))
        b, bn = plain.get_rows_sig(np.arange(total, dtype=np.uint64))
        assert np.array_equal(a, b) and np.array_equal(an.view(np.uint32), bn.view(np.uint32))
        # and against the CPU engine on the stored table
        q = a[:3]
        got = fused.query_sig(q, an[:3] if euclid else None, k)
        want = oracle.scan_batch(method, q, an[:3] if euclid else None, a, an if euclid else None, hash_num, k)[:3]
        assert_same_ranking(got, want, 3, k)
        with pytest.raises(Exception):
            fused.set_row_query_csr(total + 3, seed32, keyhash, v, k)          # skips past the end: refused, nothing stored
        assert fused.size()[0] == total
