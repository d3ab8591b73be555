"""Development sweep (not the bench): one table, many (queries_per_pass, popc_variant, nq)
settings; prints one JSON line per setting with the per-launch scan time and GB/s."""
import argparse
import json
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from jubatus_b200 import Engine, workload  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--method", default="lsh")
    ap.add_argument("--rows", type=int, default=100_000_000)
    ap.add_argument("--hash-num", type=int, default=512)
    ap.add_argument("--k", type=int, default=100)
    ap.add_argument("--qps", default="1,2,3,4,6,8,16")
    ap.add_argument("--pvs", default="0,1,2")
    ap.add_argument("--variants", default="1")
    ap.add_argument("--passes", type=int, default=6)
    ap.add_argument("--orders", default="1")
    ap.add_argument("--geoms", default="0", help="tma geometry overrides tpw:stages,...")
    a = ap.parse_args()
    dev = torch.device("cuda", 0)
    euclid = a.method == "euclid_lsh"
    w64 = (a.hash_num + 63) // 64
    eng = Engine(a.method, a.hash_num)
    eng.set_option("reserve_rows", a.rows)
    stream = torch.cuda.current_stream().cuda_stream
    gen = torch.Generator(device=dev)
    for r0 in range(0, a.rows, 4_000_000):
        n = min(4_000_000, a.rows - r0)
        gen.manual_seed(1 + r0)
        sig = torch.randint(-2 ** 63, 2 ** 63 - 1, (n, w64), dtype=torch.int64, device=dev, generator=gen)
        nrm = (torch.rand(n, device=dev, generator=gen) + 0.5) if euclid else None
        eng.set_rows_sig_dev(n, r0, sig.data_ptr(), nrm.data_ptr() if euclid else 0, stream)
        torch.cuda.synchronize()
    row_bytes = w64 * 8 + (4 if euclid else 0)
    geoms = [(0, 0)] if a.geoms == "0" else [tuple(int(y) for y in g.split(":")) for g in a.geoms.split(",")]
    for order, variant, geom in [(int(o), int(x), g) for o in a.orders.split(",") for x in a.variants.split(",") for g in geoms]:
        eng.set_option("tma_geometry", (geom[0] << 8) | geom[1])
        eng.set_option("scan_variant", variant)
        eng.set_option("scan_order", order)
        for pv in [int(x) for x in a.pvs.split(",")]:
            eng.set_option("popc_variant", pv)
            for qp in [int(x) for x in a.qps.split(",")]:
                eng.set_option("queries_per_pass", qp)
                nq = qp * a.passes
                q = torch.from_numpy(workload.random_signatures(5, nq, w64).view(np.int64)).to(dev)
                qn = torch.ones(nq, dtype=torch.float32, device=dev) if euclid else None
                oi = torch.zeros((nq, a.k), dtype=torch.int64, device=dev)
                os_ = torch.zeros((nq, a.k), dtype=torch.float32, device=dev)
                oc = torch.zeros(nq, dtype=torch.int32, device=dev)
                for rep in range(2):
                    eng.stats_reset(True)
                    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                    e0.record()
                    eng.query_sig_dev(nq, q.data_ptr(), qn.data_ptr() if euclid else 0, a.k, 0, oi.data_ptr(),
                                      os_.data_ptr(), oc.data_ptr(), stream)
                    e1.record()
                    torch.cuda.synchronize()
                st = eng.stats()
                ms = e0.elapsed_time(e1)
                avg = st["scan_ms"] / max(1, st["scan_passes"])
                print(json.dumps({"geom": "%d:%d" % geom, "order": order, "variant": variant, "pv": pv, "qp": qp, "nq": nq, "total_ms": round(ms, 3),
                                  "scan_passes": st["scan_passes"], "ms_per_pass": round(avg, 4),
                                  "scan_GBps": round(a.rows * row_bytes / (avg * 1e-3) / 1e9, 1),
                                  "qps": round(nq / (ms * 1e-3), 1),
                                  "pairs_per_s": float("%.4g" % (a.rows * qp / (avg * 1e-3)))}), flush=True)


if __name__ == "__main__":
    main()
