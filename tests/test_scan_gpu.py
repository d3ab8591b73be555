"""Parity of the CUDA scan + top-k (through the C-ABI) with the oracle: bit-exact ids, order
and scores, including ties (smaller row index first), k > N, k = 0, empty and ragged tables."""
import numpy as np
import pytest

from gpu_util import assert_same_ranking, random_rows

pytestmark = pytest.mark.gpu


def _engine(method, hash_num, **opt):
    from jubatus_b200 import Engine
    return Engine(method, hash_num, **opt)


@pytest.mark.parametrize("hash_num", [64, 100, 128, 192, 256, 512, 1024, 1])
@pytest.mark.parametrize("popc_variant,scan_variant", [(0, 1), (1, 1), (2, 1), (3, 1), (4, 1), (0, 2), (2, 2), (3, 2)])
def test_hamming_topk_random(oracle, hash_num, popc_variant, scan_variant):
    rng = np.random.RandomState(hash_num + popc_variant)
    n, nq, k = 5000, 9, 10
    rows = random_rows(rng, n, hash_num)
    q = random_rows(rng, nq, hash_num)
    q[0] = rows[17]
    with _engine("lsh", hash_num, popc_variant=popc_variant, scan_variant=scan_variant) as e:
        e.set_rows_sig(np.arange(n), rows)
        got = e.query_sig(q, None, k)
    want = oracle.scan_batch("lsh", q, None, rows, None, hash_num, k)[:3]
    assert_same_ranking(got, want, nq, k)


@pytest.mark.parametrize("scan_variant", [1, 2])
@pytest.mark.parametrize("n,k", [(0, 5), (1, 5), (31, 40), (32, 32), (33, 1), (1000, 0), (1000, 100), (4097, 257),
                                 (70000, 100), (300, 299)])
def test_hamming_sizes_and_k(oracle, n, k, scan_variant):
    rng = np.random.RandomState(n * 7 + k)
    hash_num, nq = 256, 5
    rows = random_rows(rng, n, hash_num)
    q = random_rows(rng, nq, hash_num)
    with _engine("lsh", hash_num, scan_variant=scan_variant) as e:
        if n:
            e.set_rows_sig(np.arange(n), rows)
        got = e.query_sig(q, None, k)
        assert e.size() == (n, n)
    want = oracle.scan_batch("lsh", q, None, rows, None, hash_num, k)[:3]
    assert_same_ranking(got, want, nq, k)


@pytest.mark.parametrize("scan_variant", [1, 2])
@pytest.mark.parametrize("qp", [1, 2, 3, 8])
def test_adversarial_ties(oracle, qp, scan_variant):
    # few distinct signatures -> huge tie groups; order must be by row index
    rng = np.random.RandomState(11)
    hash_num, n, nq, k = 128, 20000, 7, 100
    base = random_rows(rng, 4, hash_num)
    rows = base[rng.randint(0, 4, size=n)]
    q = np.concatenate([base, random_rows(rng, nq - 4, hash_num)])
    with _engine("lsh", hash_num, queries_per_pass=qp, scan_variant=scan_variant) as e:
        e.set_rows_sig(np.arange(n), rows)
        got = e.query_sig(q, None, k)
    want = oracle.scan_batch("lsh", q, None, rows, None, hash_num, k)[:3]
    assert_same_ranking(got, want, nq, k)


def test_all_identical_rows(oracle):
    hash_num, n, k = 512, 50000, 100
    rows = np.tile(random_rows(np.random.RandomState(1), 1, hash_num), (n, 1))
    q = rows[:2].copy()
    q[1, 0] ^= np.uint64(0xFF)
    with _engine("lsh", hash_num) as e:
        e.set_rows_sig(np.arange(n), rows)
        idx, score, count = e.query_sig(q, None, k)
    assert list(idx[0]) == list(range(k)) and list(idx[1]) == list(range(k))
    assert (score[0] == 0).all() and (score[1] == np.float32(8) / np.float32(512)).all()


def test_descending_distances_worst_case(oracle):
    # every later row is closer than all earlier ones: the threshold path is exercised at every row
    hash_num, n, k = 512, 512, 20
    rows = np.zeros((n, 8), dtype=np.uint64)
    for i in range(n):
        nbits = n - i  # distance to the all-zero query
        for b in range(nbits):
            rows[i, b // 64] |= np.uint64(1) << np.uint64(b % 64)
    q = np.zeros((1, 8), dtype=np.uint64)
    with _engine("lsh", hash_num) as e:
        e.set_rows_sig(np.arange(n), rows)
        got = e.query_sig(q, None, k)
    want = oracle.scan_batch("lsh", q, None, rows, None, hash_num, k)[:3]
    assert_same_ranking(got, want, 1, k)


@pytest.mark.parametrize("scan_variant", [1, 2])
def test_large_k_pagination(oracle, scan_variant):
    rng = np.random.RandomState(3)
    hash_num, n, nq, k = 64, 9000, 3, 5000   # > 2048 results per pass -> several pages
    rows = random_rows(rng, n, hash_num)
    q = random_rows(rng, nq, hash_num)
    with _engine("lsh", hash_num, scan_variant=scan_variant) as e:
        e.set_rows_sig(np.arange(n), rows)
        got = e.query_sig(q, None, k)
    want = oracle.scan_batch("lsh", q, None, rows, None, hash_num, k)[:3]
    assert_same_ranking(got, want, nq, k)


def test_similar_flag_and_minhash_method(oracle):
    from jubatus_b200 import SIMILAR
    rng = np.random.RandomState(8)
    hash_num, n, nq, k = 128, 3000, 4, 50
    rows = random_rows(rng, n, hash_num)
    q = random_rows(rng, nq, hash_num)
    for method in ("lsh", "minhash"):
        with _engine(method, hash_num) as e:
            e.set_rows_sig(np.arange(n), rows)
            got = e.query_sig(q, None, k, SIMILAR)
        want = oracle.scan_batch(method, q, None, rows, None, hash_num, k, similar=True)[:3]
        assert_same_ranking(got, want, nq, k)


@pytest.mark.parametrize("scan_variant", [1, 2])
@pytest.mark.parametrize("hash_num", [64, 256, 100])
def test_euclid_scores(oracle, hash_num, scan_variant):
    from jubatus_b200 import SIMILAR
    rng = np.random.RandomState(hash_num)
    n, nq, k = 6000, 6, 100
    rows = random_rows(rng, n, hash_num)
    norms = (rng.rand(n) * 10).astype(np.float32)
    norms[5] = 0.0
    norms[6] = 0.0
    q = random_rows(rng, nq, hash_num)
    qn = (rng.rand(nq) * 10).astype(np.float32)
    q[1] = rows[100]
    qn[1] = norms[100]
    qn[2] = 0.0
    with _engine("euclid_lsh", hash_num, scan_variant=scan_variant) as e:
        e.set_rows_sig(np.arange(n), rows, norms)
        got = e.query_sig(q, qn, k)
        got_sim = e.query_sig(q, qn, k, SIMILAR)
    want = oracle.scan_batch("euclid_lsh", q, qn, rows, norms, hash_num, k)[:3]
    want_sim = oracle.scan_batch("euclid_lsh", q, qn, rows, norms, hash_num, k, similar=True)[:3]
    # ids/order bit-exact; scores: same float operations -> bit-exact as well (NaN-safe compare)
    for g, w_ in ((got, want), (got_sim, want_sim)):
        for qi in range(nq):
            c = int(w_[2][qi])
            assert int(g[2][qi]) == c
            assert np.array_equal(g[0][qi, :c], w_[0][qi, :c])
            assert np.array_equal(g[1][qi, :c].view(np.uint32), w_[1][qi, :c].view(np.uint32))


def test_query_rows_uses_stored_signature(oracle):
    rng = np.random.RandomState(21)
    hash_num, n, k = 256, 4000, 10
    rows = random_rows(rng, n, hash_num)
    rows[2000] = rows[3]
    with _engine("lsh", hash_num) as e:
        e.set_rows_sig(np.arange(n), rows)
        got = e.query_rows([3, 2000, 77], k)
    want = oracle.scan_batch("lsh", rows[[3, 2000, 77]], None, rows, None, hash_num, k)[:3]
    assert_same_ranking(got, want, 3, k)
    assert list(got[0][0][:2]) == [3, 2000] and list(got[0][1][:2]) == [3, 2000]   # itself is not excluded


def test_many_queries_batch(oracle):
    rng = np.random.RandomState(4)
    hash_num, n, nq, k = 512, 20000, 300, 10
    rows = random_rows(rng, n, hash_num)
    q = random_rows(rng, nq, hash_num)
    q[:50] = rows[rng.randint(0, n, size=50)]
    with _engine("lsh", hash_num) as e:
        e.set_rows_sig(np.arange(n), rows)
        got = e.query_sig(q, None, k)
        st = e.stats()
    assert st["kernel_launches"] > 0 and st["scan_launches"] > 0
    want = oracle.scan_batch("lsh", q, None, rows, None, hash_num, k, threads=4)[:3]
    assert_same_ranking(got, want, nq, k)


def test_table_mutation_matches_column_table(oracle):
    # update in place, delete = move last row into the hole, then re-query
    rng = np.random.RandomState(6)
    hash_num, n, k = 128, 1000, 30
    rows = random_rows(rng, n, hash_num)
    with _engine("lsh", hash_num) as e:
        e.set_rows_sig(np.arange(n), rows)
        rows[10] = rows[500]
        e.set_rows_sig([10], rows[10:11])                 # update in place
        e.move_row(n - 1, 20); e.truncate(n - 1)          # delete_row(key of row 20)
        rows[20] = rows[n - 1]; rows = rows[: n - 1]
        got = e.query_sig(rows[500:502], None, k)
        back, _ = e.get_rows_sig([10, 20, 998])
    assert np.array_equal(back, rows[[10, 20, 998]])
    want = oracle.scan_batch("lsh", rows[500:502], None, rows, None, hash_num, k)[:3]
    assert_same_ranking(got, want, 2, k)


def test_errors(oracle):
    from jubatus_b200 import JubannError
    with _engine("lsh", 64) as e:
        with pytest.raises(JubannError):
            e.set_rows_sig([5], np.zeros((1, 1), dtype=np.uint64))   # skips past the end
        with pytest.raises(JubannError):
            e.query_rows([0], 1)                                       # unknown row
        with pytest.raises(JubannError):
            e.set_option("nope", 1)
    with _engine("euclid_lsh", 64) as e:
        e.set_rows_sig([0], np.zeros((1, 1), dtype=np.uint64), np.ones(1, dtype=np.float32))
        with pytest.raises(JubannError):
            e.query_sig(np.zeros((1, 1), dtype=np.uint64), None, 1)   # norms required


@pytest.mark.parametrize("scan_variant", [1, 2])
@pytest.mark.parametrize("scan_order", [0, 1])
def test_scan_order_option(oracle, scan_variant, scan_order):
    rng = np.random.RandomState(scan_order)
    hash_num, n, nq, k = 512, 150001, 5, 100
    rows = random_rows(rng, n, hash_num)
    q = random_rows(rng, nq, hash_num)
    q[0] = rows[n - 1]
    with _engine("lsh", hash_num, scan_variant=scan_variant, scan_order=scan_order) as e:
        e.set_rows_sig(np.arange(n), rows)
        got = e.query_sig(q, None, k)
    want = oracle.scan_batch("lsh", q, None, rows, None, hash_num, k, threads=4)[:3]
    assert_same_ranking(got, want, nq, k)


@pytest.mark.parametrize("scan_variant", [1, 2])
@pytest.mark.parametrize("method,hash_num,n,k", [("euclid_lsh", 256, 40000, 100), ("lsh", 64, 100000, 600),
                                                 ("lsh", 512, 70001, 1), ("euclid_lsh", 64, 140000, 2048)])
def test_sample_prepass_bound_is_safe(oracle, method, hash_num, n, k, scan_variant):
    # tables large enough for the sample pre-pass that seeds the thresholds; heavy ties in the sample
    rng = np.random.RandomState(n + k)
    nq = 6
    rows = random_rows(rng, n, hash_num)
    rows[:3000] = rows[0]                      # the sampled prefix is one repeated row
    q = random_rows(rng, nq, hash_num)
    q[0] = rows[0]
    q[1] = rows[n - 5]
    norms = (rng.rand(n) * 3).astype(np.float32) if method == "euclid_lsh" else None
    qn = (rng.rand(nq) * 3).astype(np.float32) if method == "euclid_lsh" else None
    with _engine(method, hash_num, scan_variant=scan_variant) as e:
        e.set_rows_sig(np.arange(n), rows, norms)
        got = e.query_sig(q, qn, k)
    want = oracle.scan_batch(method, q, qn, rows, norms, hash_num, k, threads=4)[:3]
    assert_same_ranking(got, want, nq, k)


@pytest.mark.parametrize("scan_variant", [1, 2])
@pytest.mark.parametrize("qp", [1, 2, 4, 16])
@pytest.mark.parametrize("hash_num,n,k", [(256, 150000, 100), (64, 90000, 7), (512, 70000, 300)])
def test_euclid_norm_bound_filter_adversarial(oracle, hash_num, n, k, qp, scan_variant):
    """euclid_lsh rows are pre-filtered by a float bound on the row norm per Hamming distance and by a
    GPU-wide histogram of score keys (tables >= 65536 rows); both may only ever let too much
    through.  Norms over six decades, zero norms, blocks of equal norms (ties by row index),
    near-copies of the queries (negative scores) and thresholds that cross zero."""
    from jubatus_b200 import SIMILAR
    rng = np.random.RandomState(hash_num * 7 + k + qp)
    nq = 16
    rows = random_rows(rng, n, hash_num)
    q = random_rows(rng, nq, hash_num)
    norms = np.exp(rng.uniform(np.log(1e-3), np.log(1e3), n)).astype(np.float32)
    norms[rng.randint(0, n, n // 50)] = 0.0
    norms[1000:3000] = np.float32(0.75)
    qn = np.exp(rng.uniform(np.log(1e-2), np.log(1e2), nq)).astype(np.float32)
    qn[3] = 0.0
    w = rows.shape[1]
    for j in range(nq // 2):      # plant near-copies with norms around the query's: negative scores
        for c in range(40):
            r = rng.randint(0, n)
            rows[r] = q[j]
            flip = rng.randint(0, hash_num, rng.randint(0, 6))
            for b in flip:
                rows[r, b // 64] ^= np.uint64(1) << np.uint64(b % 64)
            norms[r] = np.float32(qn[j] * rng.uniform(0.3, 2.5))
    assert rows.shape[1] == w
    with _engine("euclid_lsh", hash_num, scan_variant=scan_variant, queries_per_pass=qp) as e:
        e.set_rows_sig(np.arange(n), rows, norms)
        got = e.query_sig(q, qn, k)
        got_sim = e.query_sig(q, qn, k, SIMILAR)
    want = oracle.scan_batch("euclid_lsh", q, qn, rows, norms, hash_num, k, threads=4)[:3]
    assert_same_ranking(got, want, nq, k)
    want_sim = oracle.scan_batch("euclid_lsh", q, qn, rows, norms, hash_num, k, similar=True, threads=4)[:3]
    assert_same_ranking(got_sim, want_sim, nq, k)


@pytest.mark.parametrize("hash_num,n,k,qp,scan_variant", [(256, 120000, 100, 2, 0), (256, 70000, 10, 0, 1), (64, 30000, 50, 3, 0)])
def test_euclid_score_in_float_variant(oracle, hash_num, n, k, qp, scan_variant):
    """SURVEY A.9 alternate: cosf(theta) and the whole score in float (option "euclid_score" 1) against the
    oracle under the same switch -- bit-exact, on the adversarial table of the norm-bound test (the float
    bound has to allow for the float score's own rounding)."""
    import oracle_lib
    rng = np.random.RandomState(hash_num + k)
    nq = 12
    rows = random_rows(rng, n, hash_num)
    q = random_rows(rng, nq, hash_num)
    norms = np.exp(rng.uniform(np.log(1e-3), np.log(1e3), n)).astype(np.float32)
    norms[rng.randint(0, n, n // 50)] = 0.0
    norms[1000:3000] = np.float32(0.75)
    qn = np.exp(rng.uniform(np.log(1e-2), np.log(1e2), nq)).astype(np.float32)
    qn[3] = 0.0
    for j in range(nq // 2):
        for c in range(40):
            r = rng.randint(0, n)
            rows[r] = q[j]
            for b in rng.randint(0, hash_num, rng.randint(0, 6)):
                rows[r, b // 64] ^= np.uint64(1) << np.uint64(b % 64)
            norms[r] = np.float32(qn[j] * rng.uniform(0.3, 2.5))
    opts = {"queries_per_pass": qp} if qp else {}
    with _engine("euclid_lsh", hash_num, scan_variant=scan_variant, euclid_score=1, **opts) as e:
        e.set_rows_sig(np.arange(n), rows, norms)
        got = e.query_sig(q, qn, k)
        e.set_option("euclid_score", 0)
        got0 = e.query_sig(q, qn, k)
    try:
        oracle_lib.set_variants(euclid_score=1)
        want = oracle.scan_batch("euclid_lsh", q, qn, rows, norms, hash_num, k, threads=4)[:3]
    finally:
        oracle_lib.set_variants(euclid_score=0)
    assert_same_ranking(got, want, nq, k)
    want0 = oracle.scan_batch("euclid_lsh", q, qn, rows, norms, hash_num, k, threads=4)[:3]
    assert_same_ranking(got0, want0, nq, k)        # and back: the frozen choice again
    assert not np.array_equal(got[1].view(np.uint32), got0[1].view(np.uint32))
