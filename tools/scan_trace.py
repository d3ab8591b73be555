"""Tuning aid: per-group timeline of the persistent TMA scan kernel (JUBANN_SCAN_TRACE).
Needs a library built with the trace points compiled in:
    JUBANN_NVCC_EXTRA=-DJBN_SCAN_TRACE python -c "from jubatus_b200 import build; build.build_cuda()"
(touch csrc/scan.cu first; the default build leaves them out, they cost the stage loop instructions).
    python tools/scan_trace.py [method hash_num rows]
Prints, per query group, microseconds between the trace points of CTA 0 (warp 0):
prologue | stage 0 | stages 1-3 | stages 4-11 | remaining stages | final flush + barrier | write lists."""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
path = "/tmp/jubann_scan_trace.txt"
os.environ["JUBANN_SCAN_TRACE"] = path
from jubatus_b200 import Engine  # noqa: E402

method = sys.argv[1] if len(sys.argv) > 1 else "euclid_lsh"
hash_num = int(sys.argv[2]) if len(sys.argv) > 2 else 256
rows = int(sys.argv[3]) if len(sys.argv) > 3 else 12_500_000
nq, k = 64, 100
dev = torch.device("cuda:0")
w64 = (hash_num + 63) // 64
eng = Engine(method, hash_num, 0, queries_per_pass=2)
eng.set_option("reserve_rows", rows)
gen = torch.Generator(device=dev); gen.manual_seed(1)
for r0 in range(0, rows, 4_000_000):
    n = min(4_000_000, rows - r0)
    sig = torch.randint(-2 ** 63, 2 ** 63 - 1, (n, w64), dtype=torch.int64, device=dev, generator=gen)
    nrm = (torch.rand(n, device=dev, generator=gen) * 0.6 + 15.0) if method == "euclid_lsh" else None
    eng.set_rows_sig_dev(n, r0, sig.data_ptr(), nrm.data_ptr() if nrm is not None else 0, 0)
    torch.cuda.synchronize()
q = torch.randint(-2 ** 63, 2 ** 63 - 1, (nq, w64), dtype=torch.int64, device=dev, generator=gen)
qn = torch.rand(nq, device=dev, generator=gen) * 0.6 + 15.0
oi = torch.zeros((nq, k), dtype=torch.int64, device=dev); os_ = torch.zeros((nq, k), dtype=torch.float32, device=dev)
oc = torch.zeros(nq, dtype=torch.int32, device=dev)
for _ in range(3):
    eng.query_sig_dev(nq, q.data_ptr(), qn.data_ptr() if method == "euclid_lsh" else 0, k, 0, oi.data_ptr(), os_.data_ptr(), oc.data_ptr(), 0)
    torch.cuda.synchronize()
t = np.loadtxt(path, dtype=np.int64)
names = ["prologue", "stage0", "st1-3", "st4-11", "rest", "flush+bar", "write"]
for c in range(2):
    tc = t[t[:, 0] == c][:, 2:]
    d = np.diff(tc, axis=1) / 1000.0
    total = (tc[1:, 0] - tc[:-1, 0]) / 1000.0
    print("CTA %d: group period median %.1f us" % (c, np.median(total)))
    print("   " + "  ".join("%s %.1f" % (n, v) for n, v in zip(names, np.median(d, axis=0))))
    print("   p90 " + "  ".join("%s %.1f" % (n, v) for n, v in zip(names, np.percentile(d, 90, axis=0))))
