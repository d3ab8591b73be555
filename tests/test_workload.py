"""The numpy workload generator (product side) and the oracle's generator agree."""
import numpy as np

from jubatus_b200 import workload as w


def test_synth_rows_match_oracle(oracle):
    ids, vals = w.synth_rows(w.MASTER_SEED, 123, 500)
    ids2, vals2 = oracle.synth_rows(w.MASTER_SEED, 123, 500)
    assert np.array_equal(ids, ids2) and np.array_equal(vals, vals2)
    assert (np.diff(ids.astype(np.int64), axis=1) > 0).all()       # distinct, ascending
    assert (vals >= 1.0).all() and (vals < 2.0).all()


def test_feature_tables_match_oracle(oracle, feature_tables):
    s32, k64 = feature_tables
    s32o, k64o = oracle.synth_feature_tables(1 << 20)
    assert np.array_equal(s32, s32o) and np.array_equal(k64, k64o)
    assert int(k64[42]) == oracle.fnv1_64(b"42@num")


def test_random_words_match_oracle(oracle):
    assert np.array_equal(w.random_signatures(5, 100, 8).reshape(-1), oracle.fill_random_words(5, 800))


def test_torch_generator_equals_the_numpy_one():
    # bench.py builds its tables with the torch form (on the GPU); same rows, bit for bit
    import torch
    rows = torch.tensor([0, 1, 2, 5, 1000003, 99999999, 12345678] + list(range(700, 1400, 7)), dtype=torch.int64)
    for seed in (w.MASTER_SEED, w.QUERY_SEED):
        ids, vals = w.synth_rows_torch(seed, rows)
        for i, r in enumerate(rows.tolist()):
            a, b = w.synth_rows(seed, r, 1)
            assert np.array_equal(a[0].astype(np.int64), ids[i].numpy()), r
            assert np.array_equal(b[0].view(np.uint32), vals[i].numpy().view(np.uint32)), r
    ids, vals = w.synth_rows_torch(w.MASTER_SEED, torch.arange(50, dtype=torch.int64), nnz=10, nfeat_log2=4)   # forced repeats
    a, b = w.synth_rows(w.MASTER_SEED, 0, 50, nnz=10, nfeat_log2=4)
    assert np.array_equal(a.astype(np.int64), ids.numpy()) and np.array_equal(b, vals.numpy())


def test_lazy_queries_equal_make_queries():
    a, b = w.synth_rows(w.MASTER_SEED, 0, 3000)
    q1 = w.make_queries(w.MASTER_SEED, 300, a, b)
    q2 = w.make_queries_lazy(w.MASTER_SEED, 300, 3000)
    assert np.array_equal(q1[0], q2[0]) and np.array_equal(q1[1], q2[1])
    assert (q2[2][:3] >= 0).all() and (q2[3][3:6] >= 0).all() and (q2[2][3:] == -1).all()
