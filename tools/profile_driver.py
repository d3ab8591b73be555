"""Small driver for ncu captures (one case per run, so that `ncu -k regex:<kernel> -c 1` replays little):
    python tools/profile_driver.py CASE
CASE: headline | throughput | c3_hbm | c3_throughput | c4 | c5 | sig_lsh | sig_minhash
Tables are random signature words (scan cases) or synthetic datums (signature cases); 2*10^7 rows
(>> L2) unless the case says otherwise."""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from jubatus_b200 import Engine, workload  # noqa: E402

CASES = {
    # name: (method, hash_num, rows, nq, k, queries_per_pass)
    "headline": ("lsh", 512, 20_000_000, 16, 100, 2),
    "throughput": ("lsh", 512, 20_000_000, 64, 100, 0),
    "c2": ("lsh", 512, 10_000_000, 256, 10, 0),
    "c3_hbm": ("euclid_lsh", 256, 20_000_000, 16, 100, 2),
    "c3_throughput": ("euclid_lsh", 256, 20_000_000, 48, 100, 0),
    "c4": ("minhash", 128, 20_000_000, 64, 50, 0),
    "c5": ("lsh", 256, 1_000_000, 31, 10, 0),
}


def scan_case(name):
    method, hash_num, rows, nq, k, qp = CASES[name]
    dev = torch.device("cuda:0")
    w64 = (hash_num + 63) // 64
    opts = {"queries_per_pass": qp} if qp else {}
    eng = Engine(method, hash_num, 0, **opts)
    eng.set_option("reserve_rows", rows)
    gen = torch.Generator(device=dev); gen.manual_seed(1)
    euclid = method == "euclid_lsh"
    for r0 in range(0, rows, 4_000_000):
        n = min(4_000_000, rows - r0)
        sig = torch.randint(-2 ** 63, 2 ** 63 - 1, (n, w64), dtype=torch.int64, device=dev, generator=gen)
        nrm = (torch.rand(n, device=dev, generator=gen) * 0.6 + 15.0) if euclid else None
        eng.set_rows_sig_dev(n, r0, sig.data_ptr(), nrm.data_ptr() if euclid else 0, 0)
        torch.cuda.synchronize()
    q = torch.randint(-2 ** 63, 2 ** 63 - 1, (nq, w64), dtype=torch.int64, device=dev, generator=gen)
    qn = torch.rand(nq, device=dev, generator=gen) * 0.6 + 15.0
    oi = torch.zeros((nq, k), dtype=torch.int64, device=dev); os_ = torch.zeros((nq, k), dtype=torch.float32, device=dev)
    oc = torch.zeros(nq, dtype=torch.int32, device=dev)
    for _ in range(3):
        eng.query_sig_dev(nq, q.data_ptr(), qn.data_ptr() if euclid else 0, k, 0, oi.data_ptr(), os_.data_ptr(), oc.data_ptr(), 0)
        torch.cuda.synchronize()
    print(name, "ok", int(oc.min().item()))


def sig_case(method, hash_num):
    dev = torch.device("cuda:0")
    s32, k64 = workload.feature_tables(20)
    d_s32 = torch.from_numpy(s32.view(np.int32)).to(dev); d_k64 = torch.from_numpy(k64.view(np.int64)).to(dev)
    n = 500_000
    eng = Engine(method, hash_num, 0)
    indptr = torch.arange(n + 1, device=dev, dtype=torch.int64) * 100
    for rep in range(3):   # the first batch fills the projection cache (cache_insert / cache_fill), the others hit it
        ids, vals = workload.synth_rows_torch(workload.MASTER_SEED, torch.arange(rep * n, (rep + 1) * n, device=dev, dtype=torch.int64))
        flat = ids.reshape(-1)
        seed = d_s32[flat].contiguous(); key = d_k64[flat].contiguous(); v = vals.reshape(-1).contiguous()
        eng.set_rows_csr_dev(n, rep * n, n * 100, indptr.data_ptr(), 0 if method == "minhash" else seed.data_ptr(),
                             key.data_ptr() if method == "minhash" else 0, v.data_ptr(), 0)
        torch.cuda.synchronize()
    print(method, "ok", eng.size())


if __name__ == "__main__":
    case = sys.argv[1]
    if case == "sig_lsh":
        sig_case("lsh", 512)
    elif case == "sig_minhash":
        sig_case("minhash", 128)
    else:
        scan_case(case)
