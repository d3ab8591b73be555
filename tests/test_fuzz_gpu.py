"""Seeded random configurations of the scan path against the oracle: table sizes around every
planning threshold (sample pre-pass from 1024 rows, finer chunks and the LDG kernel below 1.5e5
rows, the shared histogram from 65536 rows), k around the list-reduction variants (candidate
sets of <= 512, 2048, 4096, 8192, 16384, 24576 keys and beyond), every signature width, both
kernels, 1..40 queries per pass.  Bit-exact ids, order and scores."""
import numpy as np
import pytest

from gpu_util import assert_same_ranking, random_rows

pytestmark = pytest.mark.gpu

SIZES = [1, 31, 33, 700, 1023, 1025, 4095, 4100, 33000, 65535, 65537, 150000, 152000, 310000]
HASHES = [1, 63, 64, 65, 128, 200, 256, 512, 1000, 1024]
KS = [1, 2, 3, 7, 10, 31, 33, 64, 100, 129, 170, 255, 500, 1025, 2048]


def _cases():
    # JUBANN_FUZZ_SEED / JUBANN_FUZZ_CASES widen the hunt for a one-off run; the defaults are what CI runs
    import os
    rng = np.random.RandomState(int(os.environ.get("JUBANN_FUZZ_SEED", "20261107")))
    out = []
    for i in range(int(os.environ.get("JUBANN_FUZZ_CASES", "160"))):
        method = ["lsh", "euclid_lsh", "minhash"][rng.randint(3)]
        n = SIZES[rng.randint(len(SIZES))]
        h = HASHES[rng.randint(len(HASHES))]
        k = KS[rng.randint(len(KS))]
        if n > 100000 and (k > 300 or h > 512):     # keep the oracle's share of the run short
            k = min(k, 100); h = min(h, 512)
        nq = int(rng.randint(1, 9))
        variant = int(rng.randint(0, 3))
        qp = [0, 1, 2, 3, 4, 5, 8, 16, 40][rng.randint(9)]
        out.append((i, method, n, h, k, nq, variant, qp))
    return out


@pytest.mark.parametrize("case", _cases(), ids=lambda c: "%d-%s-n%d-h%d-k%d-q%d-v%d-p%d" % c)
def test_random_configuration(oracle, case):
    from jubatus_b200 import Engine, SIMILAR
    i, method, n, h, k, nq, variant, qp = case
    rng = np.random.RandomState(1000 + i)
    rows = random_rows(rng, n, h)
    q = random_rows(rng, nq, h)
    if n > 3:                                   # ties and exact hits
        rows[n // 2] = rows[0]
        q[0] = rows[n - 1]
        rows[rng.randint(0, n, max(1, n // 50))] = q[nq - 1]
    norms = qn = None
    if method == "euclid_lsh":
        norms = np.exp(rng.uniform(np.log(0.05), np.log(20.0), n)).astype(np.float32)
        norms[rng.randint(0, n, max(1, n // 40))] = np.float32(1.5)
        qn = np.exp(rng.uniform(np.log(0.05), np.log(20.0), nq)).astype(np.float32)
    opts = {"scan_variant": variant}
    if qp:
        opts["queries_per_pass"] = qp
    flags = SIMILAR if i % 3 == 0 else 0
    with Engine(method, h, **opts) as e:
        e.set_rows_sig(np.arange(n), rows, norms)
        got = e.query_sig(q, qn, k, flags)
        got_rows = e.query_rows(np.array([0, n - 1], dtype=np.uint64), k, flags)
    want = oracle.scan_batch(method, q, qn, rows, norms, h, k, similar=bool(flags), threads=4)[:3]
    assert_same_ranking(got, want, nq, k)
    q2 = rows[[0, n - 1]]
    qn2 = norms[[0, n - 1]] if norms is not None else None
    want_rows = oracle.scan_batch(method, q2, qn2, rows, norms, h, k, similar=bool(flags), threads=4)[:3]
    assert_same_ranking(got_rows, want_rows, 2, k)
