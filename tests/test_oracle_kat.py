"""Pins of the CPU oracle against published known answers and independent restatements.

The reference's own tests hold no golden vectors for this path, so these
are the only external pins: MT19937 (init_genrand(5489)), res53 (cross-checked against
numpy's RandomState, an independent MT19937 + genrand_res53), FNV-1 64 (published test
vectors), and a pure-Python transcription of lookup8.c's mix64.
"""
import math

import numpy as np
import pytest


def test_fnv1_64_published_vectors(oracle):
    # http://www.isthe.com/chongo/tech/comp/fnv/ test vectors (FNV-1, 64 bit)
    assert oracle.fnv1_64(b"") == 0xCBF29CE484222325
    assert oracle.fnv1_64(b"a") == 0xAF63BD4C8601B7BE
    assert oracle.fnv1_64(b"foobar") == 0x340D8765A4DDA9C2


def test_fnv1_sign_extension(oracle):
    from jubatus_b200 import workload
    s = "あ@str$x".encode("utf-8")  # bytes >= 0x80 sign-extend as `char`
    assert oracle.fnv1_64(s) == workload.fnv1_64(s)


def test_mt19937_known_answers(oracle):
    out = oracle.mt_ints(5489, 10000)
    assert int(out[0]) == 3499211612          # first output of init_genrand(5489)
    assert int(out[9999]) == 4123659995       # 10000th (C++11 [rand.predef])


@pytest.mark.parametrize("seed", [0, 1, 1234, 0xFFFFFFFF, 0x9E3779B9])
def test_res53_matches_numpy(oracle, seed):
    ours = oracle.mt_doubles(seed, 1000)
    theirs = np.random.RandomState(seed).random_sample(1000)
    assert np.array_equal(ours, theirs)


def test_gaussian_is_box_muller_of_res53(oracle):
    # frozen form: sqrt(-2 log(1-a)) cos(2 pi b), a drawn first, 4 draws each
    d = oracle.mt_doubles(77, 40)
    g = oracle.mt_gaussians(77, 20)
    for j in range(20):
        a, b = d[2 * j], d[2 * j + 1]
        assert g[j] == math.sqrt(-2.0 * math.log(1.0 - a)) * math.cos(2.0 * math.pi * b)


def _mix64_py(a, b, c):
    M = (1 << 64) - 1
    for (s1, s2, s3) in ((43, 9, 8), (38, 23, 5), (35, 49, 11), (12, 18, 22)):
        a = (a - b - c) & M; a ^= c >> s1
        b = (b - c - a) & M; b ^= (a << s2) & M
        c = (c - a - b) & M; c ^= b >> s3
    return a, b, c


def test_mix64_matches_lookup8_transcription(oracle):
    rng = np.random.RandomState(3)
    for _ in range(200):
        a, b, c = (int(x) for x in rng.randint(0, 2 ** 63, size=3, dtype=np.int64).astype(np.uint64) * np.uint64(2) + np.uint64(1))
        assert oracle.mix64(a, b, c) == _mix64_py(a, b, c)


def test_minhash_calc_hash(oracle):
    lib = oracle.load()
    a, b, v = 0x0123456789ABCDEF, 17, 1.5
    c = 0xC3A5C85C97CB3127
    x, y, z = _mix64_py(a, b, c)
    x, y, z = _mix64_py(x, y, z)
    r = np.float32(x) / np.float32(0xFFFFFFFFFFFFFFFF)
    expect = -np.log(r, dtype=np.float32) / np.float32(v)
    got = lib.oracle_minhash_calc_hash(a, b, v)
    assert abs(got - float(expect)) <= 1e-6 * abs(float(expect))


def test_ranking_matches_bruteforce(oracle):
    rng = np.random.RandomState(5)
    for hash_num in (64, 100, 512):
        w = (hash_num + 63) // 64
        rows = rng.randint(0, 2 ** 63, size=(300, w), dtype=np.int64).astype(np.uint64)
        rows[50:80] = rows[10]  # ties
        q = rows[10:12].copy()
        for threads in (1, 3):
            idx, score, count, _ = oracle.scan_batch("lsh", q, None, rows, None, hash_num, 40, threads=threads)
            for qi in range(2):
                d = np.array([sum(bin(int(x)).count("1") for x in (rows[i] ^ q[qi])) for i in range(300)])
                order = sorted(range(300), key=lambda i: (d[i], i))[:40]
                assert list(idx[qi]) == order
                assert np.array_equal(score[qi], (d[order].astype(np.float32) / np.float32(hash_num)))
                assert count[qi] == 40


def test_ranking_edge_cases(oracle):
    rows = np.arange(5, dtype=np.uint64).reshape(5, 1)
    q = np.zeros((1, 1), dtype=np.uint64)
    idx, score, count, _ = oracle.scan_batch("lsh", q, None, rows, None, 64, 0)
    assert count[0] == 0
    idx, score, count, _ = oracle.scan_batch("lsh", q, None, rows, None, 64, 9)
    assert count[0] == 5 and list(idx[0][:5]) == [0, 1, 2, 4, 3]
    idx, score, count, _ = oracle.scan_batch("lsh", q, None, rows[:0], None, 64, 3)
    assert count[0] == 0


def test_engine_semantics(oracle):
    nn = oracle.OracleNN("lsh", 64)
    assert nn.set_row("a", [("k1@num", 1.0), ("k2@num", 2.0)])
    nn.set_row("b", [("k1@num", 1.0), ("k2@num", 2.0)])
    nn.set_row("c", [("k3@num", -1.0)])
    r = nn.neighbor_row_from_id("b", 2)
    assert [x[0] for x in r] == ["a", "b"] and r[0][1] == 0.0      # tie -> smaller row index first
    assert nn.neighbor_row_from_id("nope", 3) == []                # unknown id -> empty, not an error
    s = nn.similar_row_from_id("a", 1)
    assert s[0][1] == 1.0
    nn.delete_row("a")                                             # last row moves into slot 0
    assert nn.get_all_rows() == ["c", "b"]
    with pytest.raises(ValueError):
        oracle.OracleNN("lsh", 0)
    with pytest.raises(ValueError):
        oracle.OracleNN("nope", 64)


def test_mix_newer_version_wins(oracle):
    a = oracle.OracleNN("lsh", 64, my_id="A")
    b = oracle.OracleNN("lsh", 64, my_id="B")
    a.set_row("x", [("f@num", 1.0)])
    a.set_row("y", [("g@num", 1.0)])
    b.set_row("z", [("h@num", 1.0)])
    a.mix_with(b)
    assert sorted(a.get_all_rows()) == ["x", "y", "z"] and sorted(b.get_all_rows()) == ["x", "y", "z"]
    assert np.array_equal(a.get_signature("z")[0], b.get_signature("z")[0])
    b.set_row("x", [("other@num", 1.0)])   # clock on B has moved past A's version of x
    a.mix_with(b)
    assert np.array_equal(a.get_signature("x")[0], b.get_signature("x")[0])
