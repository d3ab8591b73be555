"""Generates tests/golden/nn_golden.json from the CPU oracle (oracle/jubann_oracle.hpp).

    python tests/golden/make_golden.py

The reference holds no golden vectors for this path and its arithmetic (jubatus_core 0.3.2)
cannot be built or imported here, so these fixtures do not pin the oracle to the
reference; they freeze the oracle's own answers.  A later edit that changes the oracle and the
CUDA path together -- which no oracle-versus-CUDA test can see -- fails against this file.
Contents, per engine: signatures and norms of 8 synthetic datums (10 features of 64), and the
top 10 of 3 queries over a 300-row table built from further datums (ids, float scores as bit
patterns; neighbor and similar flavours)."""
import json
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
import oracle_lib  # noqa: E402
from jubatus_b200 import workload as w  # noqa: E402

ENGINES = [("lsh", 64), ("lsh", 512), ("lsh", 100), ("euclid_lsh", 64), ("euclid_lsh", 256), ("minhash", 64), ("minhash", 128)]


def build(method, hash_num):
    s32, k64 = w.feature_tables(20)
    ids, vals = w.synth_rows(w.MASTER_SEED, 77000, 300, nnz=10, nfeat_log2=6)
    ids[150], vals[150] = ids[3], vals[3]                       # an exact duplicate: a tie by row index
    indptr, seed32, keyhash, v = w.to_csr(ids, vals, s32, k64)
    sigs, norms = oracle_lib.signatures_csr(method, hash_num, indptr, seed32, keyhash, v, threads=1)
    qids, qvals = w.synth_rows(w.MASTER_SEED ^ 0x51, 9, 3, nnz=10, nfeat_log2=6)
    qids[0], qvals[0] = ids[3], vals[3]
    qip, qs32, qk64, qv = w.to_csr(qids, qvals, s32, k64)
    qsig, qnorm = oracle_lib.signatures_csr(method, hash_num, qip, qs32, qk64, qv, threads=1)
    eu = method == "euclid_lsh"
    out = {"method": method, "hash_num": hash_num,
           "datum_signatures": [[int(x) for x in row] for row in sigs[:8]],
           "datum_norm_bits": [int(x) for x in norms[:8].view(np.uint32)] if eu else None,
           "query_signatures": [[int(x) for x in row] for row in qsig]}
    for name, sim in (("neighbor", False), ("similar", True)):
        idx, score, count = oracle_lib.scan_batch(method, qsig, qnorm if eu else None, sigs, norms if eu else None, hash_num, 10,
                                                  similar=sim)[:3]
        out[name] = {"ids": [[int(x) for x in idx[q, :int(count[q])]] for q in range(3)],
                     "score_bits": [[int(x) for x in score[q, :int(count[q])].view(np.uint32)] for q in range(3)]}
    return out


VARIANTS = [   # (name, oracle switches, engine options of the CUDA path, engines)
    ("gaussian_form=1", {"gaussian_form": 1}, {"gaussian_form": 1}, [("lsh", 64), ("lsh", 512), ("euclid_lsh", 256)]),
    ("euclid_score=1", {"euclid_score": 1}, {"euclid_score": 1}, [("euclid_lsh", 64), ("euclid_lsh", 256)]),
]


def main_variants():
    """The SURVEY A.9 alternates: the oracle's answers under each switch, frozen like the main file."""
    out = {"generator": "tests/golden/make_golden.py", "variants": []}
    for name, sw, opts, engines in VARIANTS:
        oracle_lib.set_variants(**sw)
        out["variants"].append({"name": name, "oracle": sw, "options": opts, "engines": [build(m, h) for m, h in engines]})
        oracle_lib.set_variants(gaussian_form=0, euclid_score=0)
    with open(os.path.join(HERE, "nn_golden_variants.json"), "w") as f:
        json.dump(out, f, separators=(",", ":"))
        f.write("\n")
    print("wrote", os.path.join(HERE, "nn_golden_variants.json"))


def main():
    main_variants()
    data = {"generator": "tests/golden/make_golden.py", "workload": "workload.synth_rows(MASTER_SEED, 77000, 300, nnz=10, nfeat_log2=6)",
            "engines": [build(m, h) for m, h in ENGINES]}
    with open(os.path.join(HERE, "nn_golden.json"), "w") as f:
        json.dump(data, f, separators=(",", ":"))
        f.write("\n")
    print("wrote", os.path.join(HERE, "nn_golden.json"))


if __name__ == "__main__":
    main()
