"""Per-call latency of the C-ABI entry points for one query / one row at a time (development
measurement, not the bench): python tools/latency_probe.py [rows] [hash_num] [scan_variant] [method]"""
import os
import sys
import time

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from jubatus_b200 import Engine, workload as w  # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 1_000_000
h = int(sys.argv[2]) if len(sys.argv) > 2 else 256
method = sys.argv[4] if len(sys.argv) > 4 else "lsh"
euclid = method == "euclid_lsh"
print("table: %d rows x %d bits (%s), k=10" % (n, h, method))
opts = {"scan_variant": int(sys.argv[3])} if len(sys.argv) > 3 else {}
if os.environ.get("PROBE_QP"):
    opts["queries_per_pass"] = int(os.environ["PROBE_QP"])
e = Engine(method, h, **opts)
sig = w.random_signatures(1, n, h // 64)
norms = (np.random.RandomState(1).rand(n) * 3 + 0.5).astype(np.float32) if euclid else None
e.set_rows_sig(np.arange(n), sig, norms)
q = w.random_signatures(2, 1, h // 64)
qn = np.array([1.3], dtype=np.float32) if euclid else None
def t(f, reps=300):
    for _ in range(20): f()
    t0 = time.perf_counter()
    for _ in range(reps): f()
    return (time.perf_counter() - t0) / reps * 1e6
print("query_sig host 1 query, k=10: %.1f us" % t(lambda: e.query_sig(q, qn, 10)))
ft = w.feature_tables()
ids, vals = w.synth_rows(w.MASTER_SEED, 5, 1, nnz=100, nfeat_log2=20)
csr = w.to_csr(ids, vals, *ft)
print("query_csr host 1 query: %.1f us" % t(lambda: e.query_csr(*csr, 10)))
st = e.stats(); print(st)
e.stats_reset(True)
for _ in range(100): e.query_sig(q, qn, 10)
st = e.stats(); print("scan ms per query", st["scan_ms"]/100, st)
# device-resident variant
dev = torch.device("cuda:0")
qd = torch.from_numpy(q.view(np.int64)).to(dev)
qnd = torch.from_numpy(qn).to(dev).data_ptr() if euclid else 0
oi = torch.zeros((1, 10), dtype=torch.int64, device=dev); os_ = torch.zeros((1, 10), dtype=torch.float32, device=dev); oc = torch.zeros(1, dtype=torch.int32, device=dev)
s = torch.cuda.current_stream().cuda_stream
def f():
    e.query_sig_dev(1, qd.data_ptr(), qnd, 10, 0, oi.data_ptr(), os_.data_ptr(), oc.data_ptr(), s); torch.cuda.synchronize()
print("query_sig_dev + sync: %.1f us" % t(f))
# GPU-side span of one call (events on the launching stream)
def span(k):
    oi = torch.zeros((1, k), dtype=torch.int64, device=dev); os2 = torch.zeros((1, k), dtype=torch.float32, device=dev)
    ts = []
    for _ in range(200):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize()
        e0.record()
        e.query_sig_dev(1, qd.data_ptr(), qnd, k, 0, oi.data_ptr(), os2.data_ptr(), oc.data_ptr(), s)
        e1.record()
        torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1) * 1e3)
    ts.sort()
    return ts[len(ts) // 2]
for k in (1, 10, 100):
    print("k=%d GPU span of one query_sig_dev: %.1f us" % (k, span(k)))
# one set_row (100-feature datum hashed on the device, overwriting row 5) and one query by stored row
ridx = np.array([5], dtype=np.uint64)
print("set_rows_csr host 1 row: %.1f us" % t(lambda: e.set_rows_csr(ridx, *csr)))
print("query_rows host 1 row, k=10: %.1f us" % t(lambda: e.query_rows(ridx, 10)))

# batches of stored rows as queries (light_lof re-scores up to 61 rows per add with one such call)
for nq, k in ((1, 30), (8, 10), (31, 10), (61, 10)):
    rows = (np.arange(nq, dtype=np.uint64) * 7919) % n
    e.stats_reset(True)
    us = t(lambda: e.query_rows(rows, k), reps=200)
    st = e.stats()
    print("query_rows %2d rows, k=%d: %.1f us per call; scan+sample kernels %.1f us, %d kernels per call" % (
        nq, k, us, st["scan_ms"] / 220 * 1e3, st["kernel_launches"] // 220))
