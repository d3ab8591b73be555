"""Helpers shared by the -m gpu parity tests (the CUDA path is driven through the C-ABI)."""
import numpy as np


def assert_same_ranking(got, want, nq, k_req):
    """bit-exact ids, order and float scores."""
    gidx, gscore, gcount = got
    widx, wscore, wcount = want
    for q in range(nq):
        c = int(wcount[q])
        assert int(gcount[q]) == c, "query %d: count %d != %d" % (q, int(gcount[q]), c)
        assert np.array_equal(gidx[q, :c], widx[q, :c]), "query %d: ids differ\n%s\n%s" % (q, gidx[q, :c], widx[q, :c])
        assert np.array_equal(gscore[q, :c].view(np.uint32), wscore[q, :c].view(np.uint32)), \
            "query %d: scores differ\n%s\n%s" % (q, gscore[q, :c], wscore[q, :c])


def random_rows(rng, n, hash_num):
    w = (hash_num + 63) // 64
    rows = rng.randint(0, 2 ** 63, size=(n, w), dtype=np.int64).astype(np.uint64) * np.uint64(2) + \
        rng.randint(0, 2, size=(n, w)).astype(np.uint64)
    rem = hash_num % 64
    if rem:
        rows[:, -1] &= np.uint64((1 << rem) - 1)  # bit_vector keeps unused high bits zero
    return rows
