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
