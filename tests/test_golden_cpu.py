"""The oracle against the committed fixtures of tests/golden/ (made by tests/golden/make_golden.py
from the oracle itself: they freeze its answers, they do not pin it to the reference)."""
import json
import os

import numpy as np

import oracle_lib
from jubatus_b200 import workload as w

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "nn_golden.json")


def load_golden():
    with open(GOLDEN) as f:
        return json.load(f)


def table_and_queries(feature_tables):
    s32, k64 = feature_tables
    ids, vals = w.synth_rows(w.MASTER_SEED, 77000, 300, nnz=10, nfeat_log2=6)
    ids[150], vals[150] = ids[3], vals[3]
    qids, qvals = w.synth_rows(w.MASTER_SEED ^ 0x51, 9, 3, nnz=10, nfeat_log2=6)
    qids[0], qvals[0] = ids[3], vals[3]
    return w.to_csr(ids, vals, s32, k64), w.to_csr(qids, qvals, s32, k64)


def test_oracle_reproduces_the_golden_fixtures(feature_tables):
    (indptr, seed32, keyhash, v), (qip, qs32, qk64, qv) = table_and_queries(feature_tables)
    data = load_golden()
    assert len(data["engines"]) == 7
    for e in data["engines"]:
        method, h = e["method"], e["hash_num"]
        eu = method == "euclid_lsh"
        sigs, norms = oracle_lib.signatures_csr(method, h, indptr, seed32, keyhash, v, threads=1)
        assert [[int(x) for x in r] for r in sigs[:8]] == e["datum_signatures"], (method, h)
        if eu:
            assert [int(x) for x in norms[:8].view(np.uint32)] == e["datum_norm_bits"]
        qsig, qnorm = oracle_lib.signatures_csr(method, h, qip, qs32, qk64, qv, threads=1)
        assert [[int(x) for x in r] for r in qsig] == e["query_signatures"]
        for name, sim in (("neighbor", False), ("similar", True)):
            idx, score, count = oracle_lib.scan_batch(method, qsig, qnorm if eu else None, sigs, norms if eu else None, h, 10,
                                                      similar=sim)[:3]
            for q in range(3):
                c = int(count[q])
                assert [int(x) for x in idx[q, :c]] == e[name]["ids"][q], (method, h, name, q)
                assert [int(x) for x in score[q, :c].view(np.uint32)] == e[name]["score_bits"][q], (method, h, name, q)
        # the planted duplicate: rows 3 and 150 tie, the smaller row index comes first
        assert e["neighbor"]["ids"][0][:2] == [3, 150]


VARIANTS = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "nn_golden_variants.json")


def load_variants():
    with open(VARIANTS) as f:
        return json.load(f)


def test_oracle_reproduces_the_variant_fixtures(feature_tables):
    # the SURVEY A.9 alternates (Box-Muller with sin; euclid score in float): one switch each in the
    # oracle, its answers under each switch frozen in nn_golden_variants.json
    (indptr, seed32, keyhash, v), (qip, qs32, qk64, qv) = table_and_queries(feature_tables)
    frozen = {(e["method"], e["hash_num"]): e for e in load_golden()["engines"]}
    try:
        for var in load_variants()["variants"]:
            oracle_lib.set_variants(**var["oracle"])
            for e in var["engines"]:
                method, h = e["method"], e["hash_num"]
                eu = method == "euclid_lsh"
                sigs, norms = oracle_lib.signatures_csr(method, h, indptr, seed32, keyhash, v, threads=1)
                assert [[int(x) for x in r] for r in sigs[:8]] == e["datum_signatures"], (var["name"], method, h)
                qsig, qnorm = oracle_lib.signatures_csr(method, h, qip, qs32, qk64, qv, threads=1)
                idx, score, count = oracle_lib.scan_batch(method, qsig, qnorm if eu else None, sigs, norms if eu else None, h, 10)[:3]
                for q in range(3):
                    c = int(count[q])
                    assert [int(x) for x in idx[q, :c]] == e["neighbor"]["ids"][q]
                    assert [int(x) for x in score[q, :c].view(np.uint32)] == e["neighbor"]["score_bits"][q]
                # the switch does change the answer (else the fixture would prove nothing)
                base = frozen.get((method, h))
                if base is not None:
                    assert (e["datum_signatures"], e["neighbor"]["score_bits"]) != (base["datum_signatures"], base["neighbor"]["score_bits"])
            oracle_lib.set_variants(gaussian_form=0, euclid_score=0)
    finally:
        oracle_lib.set_variants(gaussian_form=0, euclid_score=0)
