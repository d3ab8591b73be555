"""Parity of the CUDA signature kernels (through the C-ABI) with the oracle.

Tolerance (BASELINE.json north_star): >= 99.99 % of signature bits agree and a bit may
disagree only where |projection| < 1e-6 relative to the row's largest |projection|;
euclid norms within 1e-5 relative (they are in fact the same float operations)."""
import numpy as np
import pytest

from jubatus_b200 import workload as w

pytestmark = pytest.mark.gpu


def _bits(sigs, hash_num):
    b = np.unpackbits(sigs.view(np.uint8), axis=1, bitorder="little")
    return b[:, :hash_num]


def _check_bits(got, want, aux, hash_num, rel=1e-6, min_agree=0.9999):
    gb, wb = _bits(got, hash_num), _bits(want, hash_num)
    diff = gb != wb
    agree = 1.0 - diff.mean()
    assert agree >= min_agree, "bit agreement %.6f" % agree
    if diff.any():
        scale = np.abs(aux).max(axis=1, keepdims=True)
        bad = np.abs(aux)[diff] > rel * np.broadcast_to(scale, aux.shape)[diff]
        assert not bad.any(), "disagreeing bits with |projection| up to %g" % np.abs(aux)[diff].max()
    return agree


@pytest.mark.parametrize("method,hash_num", [("lsh", 64), ("lsh", 512), ("lsh", 100), ("euclid_lsh", 256),
                                             ("lsh", 1024)])
def test_projection_signatures(oracle, feature_tables, method, hash_num):
    from jubatus_b200 import Engine
    s32, k64 = feature_tables
    ids, vals = w.synth_rows(w.MASTER_SEED, 0, 400)
    indptr, seed32, keyhash, v = w.to_csr(ids, vals, s32, k64)
    with Engine(method, hash_num, cache_slots=1 << 16) as e:
        got, gnorm = e.hash_csr(indptr, seed32, None, v)
        got2, _ = e.hash_csr(indptr, seed32, None, v)     # second call: all cache hits
    want, wnorm, aux = oracle.signatures_csr(method, hash_num, indptr, seed32, None, v, want_aux=True)
    _check_bits(got, want, aux, hash_num)
    assert np.array_equal(got, got2)
    if method == "euclid_lsh":
        assert np.array_equal(gnorm.view(np.uint32), wnorm.view(np.uint32))
    else:
        assert (gnorm == 0).all()
    rem = hash_num % 64
    if rem:
        assert (got[:, -1] >> np.uint64(rem) == 0).all()


def test_single_feature_rows_pin_the_gaussians(oracle):
    # a datum with one feature of value 1: projection = the feature's gaussian column, so the
    # signature is the sign pattern of MT19937 + Box-Muller for that seed
    from jubatus_b200 import Engine
    hash_num = 512
    seeds = np.array([0, 1, 5489, 0xFFFFFFFF, 123456789, 42], dtype=np.uint32)
    indptr = np.arange(len(seeds) + 1, dtype=np.uint64)
    vals = np.ones(len(seeds), dtype=np.float32)
    with Engine("lsh", hash_num, cache_slots=1024) as e:
        got, _ = e.hash_csr(indptr, seeds, None, vals)
    for i, s in enumerate(seeds):
        col = oracle.projection_column(int(s), hash_num)
        want = np.packbits((col > 0).astype(np.uint8), bitorder="little").view(np.uint64)
        near_zero = np.abs(col) < 1e-7
        gb = np.unpackbits(got[i].view(np.uint8), bitorder="little")
        wb = np.unpackbits(want.view(np.uint8), bitorder="little")
        assert ((gb == wb) | near_zero).all()


def test_ragged_and_empty_rows(oracle, feature_tables):
    from jubatus_b200 import Engine
    s32, k64 = feature_tables
    rng = np.random.RandomState(2)
    lens = [0, 1, 31, 32, 33, 100, 0, 257, 2]
    indptr = np.concatenate([[0], np.cumsum(lens)]).astype(np.uint64)
    feats = rng.randint(0, 1 << 20, size=int(indptr[-1]))
    vals = (rng.randn(len(feats))).astype(np.float32)
    for method, hash_num in (("lsh", 128), ("euclid_lsh", 64), ("minhash", 64)):
        v = np.abs(vals) + np.float32(0.5) if method == "minhash" else vals
        with Engine(method, hash_num, cache_slots=4096) as e:
            got, gnorm = e.hash_csr(indptr, s32[feats], k64[feats], v)
        want, wnorm, aux = oracle.signatures_csr(method, hash_num, indptr, s32[feats], k64[feats], v, want_aux=True)
        assert (got[0] == 0).all() and (got[6] == 0).all()            # empty sfv -> all-zero signature
        if method == "minhash":
            assert (_bits(got, hash_num) != _bits(want, hash_num)).mean() <= 1e-3
        else:
            _check_bits(got, want, aux, hash_num)
        assert np.allclose(gnorm, wnorm, rtol=1e-6, atol=0)


@pytest.mark.parametrize("hash_num", [64, 128, 100])
def test_minhash_signatures(oracle, feature_tables, hash_num):
    from jubatus_b200 import Engine
    s32, k64 = feature_tables
    ids, vals = w.synth_rows(w.MASTER_SEED, 1000, 300)
    indptr, seed32, keyhash, v = w.to_csr(ids, vals, s32, k64)
    with Engine("minhash", hash_num) as e:
        got, _ = e.hash_csr(indptr, None, keyhash, v)
    want, _, mins = oracle.signatures_csr("minhash", hash_num, indptr, None, keyhash, v, want_aux=True)
    gb, wb = _bits(got, hash_num), _bits(want, hash_num)
    agree = 1.0 - (gb != wb).mean()
    # logf differs by <= 1 ulp between glibc and CUDA: only near-ties of the two smallest hash
    # values can flip the arg-min
    assert agree >= 0.9999, agree


@pytest.mark.parametrize("method,hash_num,slots", [("lsh", 64, 100), ("lsh", 512, 1024), ("euclid_lsh", 256, 1024),
                                                   ("lsh", 200, 1)])
def test_features_beyond_the_cache_are_drawn_on_the_spot(oracle, feature_tables, method, hash_num, slots):
    # [core] lsh_function::random_projection never refuses a feature (its cache only saves the redraw):
    # with far more distinct features than cache slots the signatures are still the oracle's
    from jubatus_b200 import Engine
    s32, k64 = feature_tables
    n = 1200                                   # ~1.1e5 distinct features of 2^20
    ids, vals = w.synth_rows(w.MASTER_SEED, 5000, n)
    indptr, seed32, keyhash, v = w.to_csr(ids, vals, s32, k64)
    assert len(np.unique(seed32)) > 50 * slots
    with Engine(method, hash_num, cache_slots=slots) as e:
        got, gnorm = e.hash_csr(indptr, seed32, None, v)          # bulk path: fills the cache, the rest misses
        got2, _ = e.hash_csr(indptr, seed32, None, v)             # again: same mix of hits and misses
        one = [e.hash_csr(*w.to_csr(ids[r:r + 1], vals[r:r + 1], s32, k64))[0][0] for r in (0, 7, n - 1)]
    want, wnorm, proj = oracle.signatures_csr(method, hash_num, indptr, seed32, None, v, want_aux=True)
    _check_bits(got, want, proj, hash_num)
    assert np.array_equal(got, got2)
    for r, sig in zip((0, 7, n - 1), one):                        # small-request path, one datum per call
        assert np.array_equal(sig, got[r])
    if method == "euclid_lsh":
        assert np.array_equal(gnorm.view(np.uint32), wnorm.view(np.uint32))


def test_set_row_and_query_past_the_cache_capacity(oracle, feature_tables):
    # the small-request paths (one set_row / one query per call) with a cache of 150 slots: rows keep
    # arriving, nothing is refused, the stored signatures and the answers stay the oracle's
    from jubatus_b200 import Engine
    from gpu_util import assert_same_ranking
    s32, k64 = feature_tables
    n, hash_num, k = 40, 64, 5
    ids, vals = w.synth_rows(w.MASTER_SEED, 0, n)
    with Engine("lsh", hash_num, cache_slots=150) as e:
        for r in range(n):
            e.set_rows_csr(np.array([r], dtype=np.uint64), *w.to_csr(ids[r:r + 1], vals[r:r + 1], s32, k64))
        assert e.size()[0] == n
        stored, _ = e.get_rows_sig(np.arange(n, dtype=np.uint64))
        qcsr = w.to_csr(ids[3:4], vals[3:4], s32, k64)
        qs, _ = e.hash_csr(*qcsr)
        got = e.query_csr(*qcsr, k)
    indptr, seed32, _, v = w.to_csr(ids, vals, s32, k64)
    want_sigs, _, proj = oracle.signatures_csr("lsh", hash_num, indptr, seed32, None, v, want_aux=True)
    _check_bits(stored, want_sigs, proj, hash_num)
    assert np.array_equal(qs[0], stored[3])
    want = oracle.scan_batch("lsh", qs, None, stored, None, hash_num, k)[:3]
    assert_same_ranking(got, want, 1, k)
    assert got[0][0][0] == 3 and got[1][0][0] == 0.0


def test_set_rows_csr_then_query_end_to_end(oracle, feature_tables):
    # the whole path: datums -> signatures -> table -> neighbor_row_from_datum
    from jubatus_b200 import Engine
    s32, k64 = feature_tables
    n, nq, k, hash_num = 3000, 20, 10, 64
    ids, vals = w.synth_rows(w.MASTER_SEED, 0, n)
    qids, qvals = w.make_queries(w.MASTER_SEED, nq, ids, vals)
    indptr, seed32, keyhash, v = w.to_csr(ids, vals, s32, k64)
    qindptr, qseed32, qkeyhash, qv = w.to_csr(qids, qvals, s32, k64)
    with Engine("lsh", hash_num, cache_slots=1 << 19) as e:
        e.set_rows_csr(np.arange(n), indptr, seed32, None, v)
        stored, _ = e.get_rows_sig(np.arange(n))
        qs, _ = e.hash_csr(qindptr, qseed32, None, qv)
        got = e.query_csr(qindptr, qseed32, None, qv, k)
    want_sigs, _ = oracle.signatures_csr("lsh", hash_num, indptr, seed32, None, v)
    assert (np.unpackbits((stored ^ want_sigs).view(np.uint8)).mean()) <= 1e-4
    # ranking must be bit-exact GIVEN the device's signatures
    want = oracle.scan_batch("lsh", qs, None, stored, None, hash_num, k)[:3]
    from gpu_util import assert_same_ranking
    assert_same_ranking(got, want, nq, k)
    assert got[1][0][0] == 0.0      # query 0 is an exact duplicate of a stored row


@pytest.mark.parametrize("method,hash_num", [("lsh", 512), ("euclid_lsh", 256), ("minhash", 128)])
def test_device_resident_csr_entry_points(feature_tables, method, hash_num):
    # jubann_hash_csr_dev / jubann_set_rows_csr_dev give what the host-buffer calls give
    import torch
    from jubatus_b200 import Engine
    s32, k64 = feature_tables
    n = 500
    ids, vals = w.synth_rows(w.MASTER_SEED, 77, n)
    indptr, seed32, keyhash, v = w.to_csr(ids, vals, s32, k64)
    dev = torch.device("cuda:0")
    d_ip = torch.from_numpy(indptr.view(np.int64)).to(dev)
    d_s = torch.from_numpy(seed32.view(np.int32)).to(dev)
    d_k = torch.from_numpy(keyhash.view(np.int64)).to(dev)
    d_v = torch.from_numpy(v).to(dev)
    stream = torch.cuda.current_stream().cuda_stream
    with Engine(method, hash_num, cache_slots=1 << 17) as e:
        want, wnorm = e.hash_csr(indptr, seed32, keyhash, v)
        sigs = torch.zeros((n, e.w64), dtype=torch.int64, device=dev)
        norms = torch.zeros(n, dtype=torch.float32, device=dev)
        e.hash_csr_dev(n, len(v), d_ip.data_ptr(), d_s.data_ptr() if method != "minhash" else 0,
                       d_k.data_ptr() if method == "minhash" else 0, d_v.data_ptr(), sigs.data_ptr(), norms.data_ptr(), stream)
        torch.cuda.synchronize()
        assert np.array_equal(sigs.cpu().numpy().view(np.uint64), want)
        assert np.array_equal(norms.cpu().numpy(), wnorm)
        e.set_rows_csr_dev(n, 0, len(v), d_ip.data_ptr(), d_s.data_ptr() if method != "minhash" else 0,
                           d_k.data_ptr() if method == "minhash" else 0, d_v.data_ptr(), stream)
        torch.cuda.synchronize()
        stored, snorm = e.get_rows_sig(np.arange(n))
        assert e.size() == (n, n)
        assert np.array_equal(stored, want) and np.array_equal(snorm, wnorm)


@pytest.mark.parametrize("method,hash_num", [("lsh", 512), ("euclid_lsh", 256), ("minhash", 128), ("lsh", 64)])
def test_small_request_path_equals_batch_path(feature_tables, method, hash_num):
    """Requests under 1 MiB are staged in one copy and hashed without reading the cache counters
    back; larger ones take the chunked path.  Same rows, same signatures, same answers; rows with
    no features and empty requests included."""
    from jubatus_b200 import Engine
    s32, k64 = feature_tables
    n_big = 6000                                   # 6000 x 100 features: > 1 MiB of CSR
    ids, vals = w.synth_rows(w.MASTER_SEED, 4242, n_big)
    indptr, seed32, keyhash, v = w.to_csr(ids, vals, s32, k64)
    # rows 1 and 4 of the small request have no features at all
    nnz = np.array([100, 0, 100, 37, 0, 100], dtype=np.uint64)
    sip = np.concatenate([[0], np.cumsum(nnz)]).astype(np.uint64)
    pick = np.concatenate([np.arange(r * 100, r * 100 + int(c)) for r, c in enumerate(nnz)]).astype(np.int64)
    s_seed, s_key, s_val = seed32[pick], keyhash[pick], v[pick]
    with Engine(method, hash_num) as big, Engine(method, hash_num) as small:
        # the same six rows, once inside a large batch (rows 0..5 rebuilt with the ragged lengths)
        rag_ip = np.concatenate([sip, sip[-1] + np.arange(1, n_big - 6 + 1, dtype=np.uint64) * np.uint64(100)])
        rag_pick = np.concatenate([pick, np.arange(600, n_big * 100)])
        big.set_rows_csr(np.arange(n_big, dtype=np.uint64), rag_ip, seed32[rag_pick], keyhash[rag_pick], v[rag_pick])
        small.set_rows_csr(np.arange(6, dtype=np.uint64), sip, s_seed, s_key, s_val)
        bs, bn = big.get_rows_sig(np.arange(6, dtype=np.uint64))
        ss, sn = small.get_rows_sig(np.arange(6, dtype=np.uint64))
        assert np.array_equal(bs, ss) and np.array_equal(bn.view(np.uint32), sn.view(np.uint32))
        if method != "minhash":
            assert not ss[1].any() and not ss[4].any()          # empty sfv -> all-zero signature
        # queries: six datums through the small path against the whole batch through the large one
        k = 5
        got_small = big.query_csr(sip, s_seed, s_key, s_val, k)
        q_ip = rag_ip[:n_big + 1]
        got_big = big.query_csr(q_ip, seed32[rag_pick], keyhash[rag_pick], v[rag_pick], k)
        for a, b in zip(got_small, got_big):
            assert np.array_equal(np.asarray(a)[:6], np.asarray(b)[:6])
        # an empty request is fine on both paths
        e0 = small.query_csr(np.zeros(1, dtype=np.uint64), s_seed[:0], s_key[:0], s_val[:0], k)
        assert len(e0[2]) == 0
