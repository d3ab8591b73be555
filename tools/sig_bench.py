"""Signature-kernel throughput (development measurement, not the bench): synthetic datums of 100
features out of 2^20 generated on the device, hashed with jubann_hash_csr_dev.  Reports datums/s
and the gathered GB/s (nnz x hash_num x 4 B per datum -- SURVEY 8d's figure for this kernel)."""
import json
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from jubatus_b200 import Engine, workload  # noqa: E402


def main():
    dev = torch.device("cuda", 0)
    s32, k64 = workload.feature_tables(20)
    d_s32 = torch.from_numpy(s32.view(np.int32)).to(dev)
    d_k64 = torch.from_numpy(k64.view(np.int64)).to(dev)
    nnz_per, n = 100, 1_000_000
    gen = torch.Generator(device=dev); gen.manual_seed(3)
    ids = torch.randint(0, 1 << 20, (n, nnz_per), device=dev, generator=gen)
    vals = (torch.rand((n, nnz_per), device=dev, generator=gen) + 1.0).reshape(-1).contiguous()
    seed32 = d_s32[ids.reshape(-1)].contiguous()
    key64 = d_k64[ids.reshape(-1)].contiguous()
    indptr = (torch.arange(n + 1, device=dev, dtype=torch.int64) * nnz_per).contiguous()
    stream = torch.cuda.current_stream().cuda_stream
    for method, h in (("lsh", 512), ("lsh", 64), ("lsh", 128), ("euclid_lsh", 256), ("minhash", 128)):
        with Engine(method, h) as e:
            w64 = (h + 63) // 64
            sigs = torch.zeros((n, w64), dtype=torch.int64, device=dev)
            norms = torch.zeros(n, dtype=torch.float32, device=dev)
            args = (n, n * nnz_per, indptr.data_ptr(), seed32.data_ptr() if method != "minhash" else 0,
                    key64.data_ptr() if method == "minhash" else 0, vals.data_ptr(), sigs.data_ptr(), norms.data_ptr(), stream)
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(); e.hash_csr_dev(*args); e1.record(); torch.cuda.synchronize()
            first_ms = e0.elapsed_time(e1)          # includes generating the projection cache rows
            best = 1e30
            for _ in range(3):
                e0.record(); e.hash_csr_dev(*args); e1.record(); torch.cuda.synchronize()
                best = min(best, e0.elapsed_time(e1))
            out = {"method": method, "hash_num": h, "datums": n, "nnz": nnz_per, "first_call_ms": round(first_ms, 2),
                   "steady_ms": round(best, 3), "datums_per_s": round(n / (best * 1e-3)),
                   "gathered_GBps": round(n * nnz_per * h * 4 / (best * 1e-3) / 1e9, 1) if method != "minhash" else None,
                   "hash_evals_per_s": float("%.4g" % (n * nnz_per * h / (best * 1e-3))) if method == "minhash" else None}
            print(json.dumps(out), flush=True)


if __name__ == "__main__":
    main()
