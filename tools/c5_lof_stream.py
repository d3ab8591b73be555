"""Config 5 on one GPU (development measurement, not the bench): jubaanomaly light_lof over
lsh hash_num=256, nearest_neighbor_num=10, reverse_nearest_neighbor_num=30.  The table is bulk
loaded to --rows rows, then --adds streaming `add` calls (each: set_row, two reverse-neighbor
collections, one batched re-scoring of up to 61 rows, calc_score of the new row) and --scores
`calc_score` calls are timed through the host layer.  One JSON line."""
import argparse
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from jubatus_b200 import nn, workload as w  # noqa: E402


def datum_of(ids, vals):
    d = nn.Datum()
    for f, v in zip(ids, vals):
        d.add_number(str(int(f)), float(v))
    return d


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--rows", type=int, default=1_000_000)
    ap.add_argument("--adds", type=int, default=300)
    ap.add_argument("--scores", type=int, default=300)
    ap.add_argument("--nnz", type=int, default=100)
    ap.add_argument("--method", default="lsh")
    ap.add_argument("--hash-num", type=int, default=256)
    a = ap.parse_args()
    conv = {"string_filter_types": {}, "string_filter_rules": [], "num_filter_types": {}, "num_filter_rules": [],
            "string_types": {}, "string_rules": [], "num_types": {}, "num_rules": [{"key": "*", "type": "num"}]}
    cfg = json.dumps({"method": "light_lof", "converter": conv,
                      "parameter": {"nearest_neighbor_num": 10, "reverse_nearest_neighbor_num": 30, "method": a.method,
                                    "parameter": {"hash_num": a.hash_num}}})
    an = nn.Anomaly(cfg)
    t0 = time.time()
    chunk = 50_000
    for r0 in range(0, a.rows, chunk):
        n = min(chunk, a.rows - r0)
        ids, vals = w.synth_rows(w.MASTER_SEED, r0, n, nnz=a.nnz, nfeat_log2=20)
        an.add_bulk([datum_of(ids[i], vals[i]) for i in range(n)])
    t_load = time.time() - t0
    ids, vals = w.synth_rows(w.MASTER_SEED ^ 0x51, 0, a.adds + a.scores, nnz=a.nnz, nfeat_log2=20)
    ds = [datum_of(ids[i], vals[i]) for i in range(a.adds + a.scores)]
    for d in ds[:5]:
        an.calc_score(d)
    t0 = time.time()
    scores = [an.add(d)[1] for d in ds[:a.adds]]
    t_add = time.time() - t0
    t0 = time.time()
    s2 = [an.calc_score(d) for d in ds[a.adds:]]
    t_score = time.time() - t0
    finite = [s for s in scores + s2 if s == s and s not in (float("inf"), float("-inf"))]
    print(json.dumps({"config": "light_lof over %s hash_num=%d, nn=10, rnn=30, nnz=%d" % (a.method, a.hash_num, a.nnz),
                      "rows": a.rows, "bulk_load_s": round(t_load, 2),
                      "adds": a.adds, "adds_per_s": round(a.adds / t_add, 1), "ms_per_add": round(t_add / a.adds * 1e3, 3),
                      "calc_scores": a.scores, "calc_scores_per_s": round(a.scores / t_score, 1),
                      "ms_per_calc_score": round(t_score / a.scores * 1e3, 3),
                      "mean_score": round(sum(finite) / max(1, len(finite)), 4),
                      "note": "single caller, one request at a time, through jubatus_b200.nn (ctypes) -> C++ host layer -> C-ABI"}))


if __name__ == "__main__":
    main()
