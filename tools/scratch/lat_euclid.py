import sys, time, numpy as np, torch
sys.path.insert(0, '/root/repo')
from jubatus_b200 import Engine, workload as w
n, h = 1_000_000, 256
e = Engine("euclid_lsh", h)
sig = w.random_signatures(1, n, h // 64)
rng = np.random.RandomState(1)
norms = (rng.rand(n) * 3 + 0.5).astype(np.float32)
e.set_rows_sig(np.arange(n), sig, norms)
q = w.random_signatures(2, 1, h // 64); qn = np.array([1.3], dtype=np.float32)
def t(f, reps=300):
    for _ in range(20): f()
    t0 = time.perf_counter()
    for _ in range(reps): f()
    return (time.perf_counter() - t0) / reps * 1e6
print("euclid 1M x 256: query_sig host 1 query k=10: %.1f us" % t(lambda: e.query_sig(q, qn, 10)))
print("euclid 1M x 256: query_sig host 1 query k=100: %.1f us" % t(lambda: e.query_sig(q, qn, 100)))
