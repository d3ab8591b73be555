"""The reference-shaped host API (config JSON, string ids, datums) on the GPU against the
oracle engine: these read like client_test/nearest_neighbor_test.cpp, plus the numeric
assertions the reference's own tests lack."""
import json

import numpy as np
import pytest

from jubatus_b200 import nn, workload as w

pytestmark = pytest.mark.gpu

CONVERTER = {
    "string_filter_types": {}, "string_filter_rules": [], "num_filter_types": {}, "num_filter_rules": [],
    "string_types": {}, "string_rules": [{"key": "*", "type": "str", "sample_weight": "bin", "global_weight": "bin"}],
    "num_types": {}, "num_rules": [{"key": "*", "type": "num"}],
}


def config(method, hash_num, **extra):
    c = {"method": method, "parameter": dict({"hash_num": hash_num}, **extra), "converter": CONVERTER}
    return json.dumps(c)


def datum_of(ids, vals):
    d = nn.Datum()
    for f, v in zip(ids, vals):
        d.add_number(str(int(f)), float(v))
    return d


def sfv_of(ids, vals):
    return [("%d@num" % int(f), float(v)) for f, v in zip(ids, vals)]


def same(got, want, rel=0.0):
    assert [g[0] for g in got] == [x[0] for x in want]
    for g, x in zip(got, want):
        assert g[1] == x[1] or abs(g[1] - x[1]) <= rel * abs(x[1]), (g, x)


@pytest.mark.parametrize("method,hash_num", [("lsh", 64), ("euclid_lsh", 64), ("minhash", 64), ("lsh", 512)])
def test_rpc_surface_matches_oracle(oracle, method, hash_num):
    n, k = 300, 10
    ids, vals = w.synth_rows(w.MASTER_SEED, 0, n, nnz=20, nfeat_log2=10)
    cli = nn.NearestNeighbor(config(method, hash_num))
    ora = oracle.OracleNN(method, hash_num)
    for r in range(n):
        assert cli.set_row("row%d" % r, datum_of(ids[r], vals[r]))
        ora.set_row("row%d" % r, sfv_of(ids[r], vals[r]))
    assert cli.size() == n and cli.get_all_rows() == ora.get_all_rows()
    # stored signatures: bit agreement (minhash: logf ulp differences may flip a near-tie)
    diff = sum(bin(int(a ^ b)).count("1") for r in range(0, n, 7)
               for a, b in zip(cli.get_signature("row%d" % r, hash_num)[0], ora.get_signature("row%d" % r)[0]))
    assert diff <= 1
    if diff == 0:
        for r in (0, 17, 299):
            same(cli.neighbor_row_from_id("row%d" % r, k), ora.neighbor_row_from_id("row%d" % r, k), 1e-5)
            same(cli.similar_row_from_id("row%d" % r, k), ora.similar_row_from_id("row%d" % r, k), 1e-5)
            q = datum_of(ids[r][:15], vals[r][:15])
            qs = sfv_of(ids[r][:15], vals[r][:15])
            same(cli.neighbor_row_from_datum(q, k), ora.neighbor_row_from_datum(qs, k), 1e-5)
            same(cli.similar_row_from_datum(q, k), ora.similar_row_from_datum(qs, k), 1e-5)
    assert cli.neighbor_row_from_id("no such id", 3) == []
    assert cli.clear() and cli.size() == 0 and cli.get_all_rows() == []


def test_client_test_sequence():
    # client_test/nearest_neighbor_test.cpp:57-98, with results checked
    cli = nn.NearestNeighbor(config("lsh", 64))
    d = nn.Datum().add_number("key", 1.0)
    assert cli.set_row("id", d)
    assert cli.set_row("neighbor_row_from_id", d)
    assert cli.neighbor_row_from_id("neighbor_row_from_id", 1) == [("id", 0.0)]
    assert cli.neighbor_row_from_datum(d, 5) == [("id", 0.0), ("neighbor_row_from_id", 0.0)]
    assert cli.similar_row_from_id("id", 1) == [("id", 1.0)]
    assert cli.similar_row_from_datum(d, 1) == [("id", 1.0)]
    assert json.loads(cli.get_config())["method"] == "lsh"
    assert cli.clear()


def test_update_delete_and_batched_set_rows(oracle):
    hash_num, n, k = 128, 200, 20
    ids, vals = w.synth_rows(w.MASTER_SEED, 500, n, nnz=10, nfeat_log2=8)
    cli = nn.NearestNeighbor(config("lsh", hash_num))
    ora = oracle.OracleNN("lsh", hash_num)
    cli.set_rows(["r%d" % r for r in range(n)], [datum_of(ids[r], vals[r]) for r in range(n)])
    for r in range(n):
        ora.set_row("r%d" % r, sfv_of(ids[r], vals[r]))
    cli.set_row("r5", datum_of(ids[77], vals[77])); ora.set_row("r5", sfv_of(ids[77], vals[77]))   # update in place
    for victim in ("r10", "r199", "r0"):
        assert cli.delete_row(victim) and ora.delete_row(victim)                                       # swap-delete
    assert not cli.delete_row("r10")
    assert cli.get_all_rows() == ora.get_all_rows()
    for r in ("r5", "r77", "r198"):
        same(cli.neighbor_row_from_id(r, k), ora.neighbor_row_from_id(r, k))


def test_config_errors():
    with pytest.raises(nn.JubatusError, match="unsupported_method"):
        nn.NearestNeighbor(config("nope", 64))
    with pytest.raises(nn.JubatusError, match="hash_num"):
        nn.NearestNeighbor(config("lsh", 0))
    with pytest.raises(nn.JubatusError, match="unknown parameter key"):
        nn.NearestNeighbor(config("lsh", 64, bogus=1))
    with pytest.raises(nn.JubatusError):
        nn.NearestNeighbor("{")
    nn.NearestNeighbor(config("lsh", 64, threads=2)).close()       # lsh_multithread.json is accepted


def test_mix_between_replicas(oracle):
    hash_num = 64
    ids, vals = w.synth_rows(w.MASTER_SEED, 900, 40, nnz=10, nfeat_log2=8)
    a, b = nn.NearestNeighbor(config("lsh", hash_num), my_id="A"), nn.NearestNeighbor(config("lsh", hash_num), my_id="B")
    oa, ob = oracle.OracleNN("lsh", hash_num, my_id="A"), oracle.OracleNN("lsh", hash_num, my_id="B")
    for r in range(40):
        tgt, otgt = (a, oa) if r % 2 == 0 else (b, ob)
        tgt.set_row("r%d" % r, datum_of(ids[r], vals[r])); otgt.set_row("r%d" % r, sfv_of(ids[r], vals[r]))
    a.mix_with(b); oa.mix_with(ob)
    assert a.get_all_rows() == oa.get_all_rows() and b.get_all_rows() == ob.get_all_rows()
    b.set_row("r0", datum_of(ids[1], vals[1])); ob.set_row("r0", sfv_of(ids[1], vals[1]))   # newer version on B
    a.mix_with(b); oa.mix_with(ob)
    for r in ("r0", "r1", "r39"):
        assert np.array_equal(a.get_signature(r, hash_num)[0], oa.get_signature(r)[0])
        same(a.neighbor_row_from_id(r, 10), oa.neighbor_row_from_id(r, 10))
        same(b.neighbor_row_from_id(r, 10), ob.neighbor_row_from_id(r, 10))


def test_recommender_nearest_neighbor_recommender(oracle):
    cfg = json.dumps({"method": "nearest_neighbor_recommender", "converter": CONVERTER,
                      "parameter": {"method": "minhash", "parameter": {"hash_num": 128}}})
    rec = nn.Recommender(cfg)
    ora = oracle.OracleNN("minhash", 128)
    ids, vals = w.synth_rows(w.MASTER_SEED, 40, 60, nnz=12, nfeat_log2=8)
    for r in range(60):
        rec.update_row("u%d" % r, datum_of(ids[r], vals[r])); ora.set_row("u%d" % r, sfv_of(ids[r], vals[r]))
    q = datum_of(ids[3], vals[3])
    got, want = rec.similar_row_from_datum(q, 10), ora.similar_row_from_datum(sfv_of(ids[3], vals[3]), 10)
    assert got[0] == ("u3", 1.0)
    assert [g[0] for g in got][:3] == [x[0] for x in want][:3]
    assert rec.similar_row_from_id("u3", 1) == [("u3", 1.0)]
    assert rec.similar_row_from_id("nobody", 3) == []
    rec.clear_row("u3")
    assert "u3" not in rec.get_all_rows()
    with pytest.raises(nn.JubatusError):
        nn.Recommender(json.dumps({"method": "lsh", "converter": CONVERTER, "parameter": {"method": "lsh", "parameter": {}}}))


def test_anomaly_light_lof(oracle):
    cfg = json.dumps({"method": "light_lof", "converter": CONVERTER,
                      "parameter": {"nearest_neighbor_num": 3, "reverse_nearest_neighbor_num": 6, "method": "lsh",
                                    "parameter": {"hash_num": 256}}})
    an = nn.Anomaly(cfg)
    ora = oracle.OracleNN("lsh", 256, lof=(3, 6))
    ids, vals = w.synth_rows(w.MASTER_SEED, 70, 40, nnz=8, nfeat_log2=6)
    for r in range(40):
        rid, score = an.add(datum_of(ids[r], vals[r]))
        ora.set_row(str(r), sfv_of(ids[r], vals[r]))
        want = ora.calc_score_id(str(r))
        assert rid == str(r)
        assert score == want or abs(score - want) <= 1e-5 * abs(want) or (np.isinf(score) and np.isinf(want)), (r, score, want)
    for r in (0, 7, 39):
        got, want = an.lof_params(str(r)), ora.lof_params(str(r))
        assert np.allclose(got, want, rtol=1e-5)
    q = datum_of(ids[5][:6], vals[5][:6])
    assert abs(an.calc_score(q) - ora.calc_score(sfv_of(ids[5][:6], vals[5][:6]))) <= 1e-5
    with pytest.raises(nn.JubatusError):
        nn.Anomaly(json.dumps({"method": "lof", "converter": CONVERTER, "parameter": {"method": "lsh", "parameter": {}}}))


def test_anomaly_numeric_entry_points_equal_the_datum_ones():
    # add_numeric / calc_score_numeric / add_bulk_numeric build the datum in C++ from feature ids: same
    # rows, scores and parameters as datums marshalled key by key
    cfg = json.dumps({"method": "light_lof", "converter": CONVERTER,
                      "parameter": {"nearest_neighbor_num": 3, "reverse_nearest_neighbor_num": 6, "method": "lsh",
                                    "parameter": {"hash_num": 128}}})
    a, b = nn.Anomaly(cfg), nn.Anomaly(cfg)
    ids, vals = w.synth_rows(w.MASTER_SEED, 11, 60, nnz=8, nfeat_log2=6)
    a.add_bulk([datum_of(ids[r], vals[r]) for r in range(30)])
    assert b.add_bulk_numeric(ids[:30], vals[:30]) == 0
    for r in range(30, 50):
        ra, sa = a.add(datum_of(ids[r], vals[r]))
        rb, sb = b.add_numeric(ids[r], vals[r])
        assert ra == rb and (sa == sb or (np.isnan(sa) and np.isnan(sb))), (r, sa, sb)
    for r in range(50, 60):
        sa, sb = a.calc_score(datum_of(ids[r], vals[r])), b.calc_score_numeric(ids[r], vals[r])
        assert sa == sb or (np.isnan(sa) and np.isnan(sb))
    assert a.table_digest() == b.table_digest()
    for r in (0, 29, 31, 49):
        assert a.lof_params(str(r)) == b.lof_params(str(r))
    b.add_numeric(ids[0], vals[0])
    assert a.table_digest() != b.table_digest()


def test_concurrent_queries_are_coalesced_by_the_batcher(oracle):
    # the server answers queries from several RPC worker threads without the model lock
    # (nearest_neighbor_impl.cpp:75-102); the batcher turns them into batched launches
    import threading
    hash_num, n, k = 128, 400, 10
    ids, vals = w.synth_rows(w.MASTER_SEED, 300, n, nnz=12, nfeat_log2=9)
    cli = nn.NearestNeighbor(config("lsh", hash_num))
    cli.set_rows(["r%d" % r for r in range(n)], [datum_of(ids[r], vals[r]) for r in range(n)])
    want = [cli.neighbor_row_from_datum(datum_of(ids[r][:9], vals[r][:9]), k) for r in range(64)]
    cli.enable_batching(max_batch=16, max_wait_us=20000)
    got = [None] * 64
    errors = []

    def worker(t):
        try:
            for r in range(t, 64, 8):
                got[r] = cli.neighbor_row_from_datum(datum_of(ids[r][:9], vals[r][:9]), k)
        except Exception as e:   # noqa: BLE001
            errors.append(e)

    threads = [threading.Thread(target=worker, args=(t,)) for t in range(8)]
    for t in threads:
        t.start()
    for t in threads:
        t.join()
    assert not errors, errors
    assert got == want
    st = cli.batch_stats()
    assert st["queries"] == 64 and st["batches"] < 64     # several queries shared a launch
    assert cli.similar_row_from_datum(datum_of(ids[0], vals[0]), 1)[0] == ("r0", 1.0)


def test_engine_is_thread_safe_at_the_c_abi(oracle):
    # concurrent queries + one writer on one handle: every answer must be the exact ranking of
    # the table either before or after the write (calls are serialised inside the library)
    import threading
    from jubatus_b200 import Engine
    rng = np.random.RandomState(0)
    hash_num, n, k = 256, 30000, 20
    rows = rng.randint(0, 2 ** 62, size=(n, 4), dtype=np.int64).astype(np.uint64)
    q = rows[:32].copy()
    new_row = rows[:1] ^ np.uint64(0xFFFFFFFF)
    before = oracle.scan_batch("lsh", q, None, rows, None, hash_num, k, threads=4)[:3]
    after = oracle.scan_batch("lsh", q, None, np.concatenate([rows, new_row]), None, hash_num, k, threads=4)[:3]
    with Engine("lsh", hash_num) as e:
        e.set_rows_sig(np.arange(n), rows)
        out = [[] for _ in range(8)]

        def worker(t):
            for _ in range(10):
                out[t].append(e.query_sig(q[t * 4:(t + 1) * 4], None, k))

        ts = [threading.Thread(target=worker, args=(t,)) for t in range(8)]
        for t in ts:
            t.start()
        e.set_rows_sig([n], new_row)
        for t in ts:
            t.join()
        assert e.size() == (n + 1, n + 1)
        for t in range(8):
            sl = slice(t * 4, (t + 1) * 4)
            for g in out[t]:
                ok_b = np.array_equal(g[0], before[0][sl]) and np.array_equal(g[1], before[1][sl])
                ok_a = np.array_equal(g[0], after[0][sl]) and np.array_equal(g[1], after[1][sl])
                assert ok_b or ok_a, t
        final = e.query_sig(q, None, k)
        assert np.array_equal(final[0], after[0]) and np.array_equal(final[1], after[1])


def test_pack_unpack_roundtrip_and_status(oracle):
    hash_num, n, k = 64, 120, 10
    ids, vals = w.synth_rows(w.MASTER_SEED, 1200, n, nnz=10, nfeat_log2=8)
    a = nn.NearestNeighbor(config("euclid_lsh", hash_num), my_id="A")
    for r in range(n):
        a.set_row("r%d" % r, datum_of(ids[r], vals[r]))
    a.delete_row("r3")
    st = json.loads(a.get_status()["data"])
    assert st["rows"] == n - 1 and st["hash_num"] == hash_num
    blob = a.pack()
    b = nn.NearestNeighbor(config("euclid_lsh", hash_num), my_id="B")
    b.set_row("junk", datum_of(ids[0], vals[0]))
    b.unpack(blob)
    assert b.get_all_rows() == a.get_all_rows()
    for r in ("r0", "r50", "r119"):
        assert b.neighbor_row_from_id(r, k) == a.neighbor_row_from_id(r, k)
    b.set_row("new", datum_of(ids[5], vals[5]))          # the clock continues where the saved table stopped
    assert b.size() == n
    with pytest.raises(nn.JubatusError):
        nn.NearestNeighbor(config("euclid_lsh", 128)).unpack(blob)   # config mismatch is refused, as load does
    with pytest.raises(nn.JubatusError):
        b.unpack(blob[:40])


def test_baseline_config1_lsh64_10k_rows_k10(oracle, feature_tables):
    """BASELINE.json config 1: jubanearest_neighbor method=lsh hash_num=64, 10k synthetic sparse rows
    (100 nnz out of 2^20 features), neighbor_row_from_datum k=10 -- the standalone CPU server is the
    oracle engine here (the real binary cannot be built), the GPU answers through the same RPC names."""
    n, nq, k, hash_num = 10000, 200, 10, 64
    ids, vals = w.synth_rows(w.MASTER_SEED, 0, n)
    qids, qvals = w.make_queries(w.MASTER_SEED, nq, ids, vals)
    cli = nn.NearestNeighbor(config("lsh", hash_num))
    ora = oracle.OracleNN("lsh", hash_num)
    cli.set_rows(["%d" % r for r in range(n)], [datum_of(ids[r], vals[r]) for r in range(n)])
    for r in range(n):
        ora.set_row("%d" % r, sfv_of(ids[r], vals[r]))
    # signatures: all bits equal on this data set (tolerance would allow 1e-4)
    diff = sum(bin(int(a ^ b)).count("1") for r in range(0, n, 50)
               for a, b in zip(cli.get_signature("%d" % r, hash_num)[0], ora.get_signature("%d" % r)[0]))
    assert diff == 0
    for j in range(nq):
        got = cli.neighbor_row_from_datum(datum_of(qids[j], qvals[j]), k)
        want = ora.neighbor_row_from_datum(sfv_of(qids[j], qvals[j]), k)
        assert got == want, j
    assert cli.neighbor_row_from_datum(datum_of(qids[0], qvals[0]), 1)[0][1] == 0.0   # query 0 is a stored row


def test_save_load_rpcs_write_the_reference_model_file_container(tmp_path):
    """save / load / load_file (framework/server_base.cpp:135-212) through the ".jubatus" container of
    framework/save_load.cpp; the header is re-read here independently of the library."""
    import struct
    import zlib
    hash_num, n, k = 128, 300, 10
    ids, vals = w.synth_rows(w.MASTER_SEED, 77, n, nnz=12, nfeat_log2=9)
    cfg = config("lsh", hash_num)
    a = nn.NearestNeighbor(cfg, my_id="A")
    a.set_storage(str(tmp_path), "10.0.0.7", 9199)
    for r in range(n):
        a.set_row("r%d" % r, datum_of(ids[r], vals[r]))
    want = {r: a.neighbor_row_from_id(r, k) for r in ("r0", "r17", "r299")}
    with pytest.raises(nn.JubatusError, match="empty id is not allowed"):
        a.save("")
    out = a.save("snap1")
    path = str(tmp_path / "10.0.0.7_9199_nearest_neighbor_snap1.jubatus")     # build_local_path, server_base.cpp:41-50
    assert out == {"10.0.0.7_9199": path}
    f = open(path, "rb").read()
    fmt, major, minor, maint, crc, sys_n, user_n = struct.unpack(">QIIIIQQ", f[8:48])
    assert f[:8] == b"jubatus\0" and (fmt, major, minor, maint) == (1, 0, 9, 2) and 48 + sys_n + user_n == len(f)
    assert crc == zlib.crc32(f[48:], zlib.crc32(f[32:48], zlib.crc32(f[:28])))
    # load into the same server after clear
    a.clear()
    assert a.size() == 0
    assert a.load("snap1") is True
    assert a.size() == n
    for r, res in want.items():
        assert a.neighbor_row_from_id(r, k) == res
    # a server with another config refuses the file on `load` ...
    b = nn.NearestNeighbor(config("lsh", 64), my_id="B")
    b.set_storage(str(tmp_path), "10.0.0.7", 9199)
    with pytest.raises(nn.JubatusError, match="server config mismatched"):
        b.load("snap1")
    # ... and takes the file's config with load_file (--model_file at start-up)
    b.load_file(path)
    assert json.loads(b.get_config()) == json.loads(cfg)
    assert b.size() == n
    for r, res in want.items():
        assert b.neighbor_row_from_id(r, k) == res
    b.set_row("extra", datum_of(ids[1], vals[1]))
    assert b.size() == n + 1
    # same config written differently is the same config
    c = nn.NearestNeighbor(json.dumps(json.loads(cfg), indent=4, sort_keys=True), my_id="C")
    c.set_storage(str(tmp_path), "10.0.0.7", 9199)
    assert c.load("snap1") and c.size() == n
    # damage is detected before anything is unpacked
    g = bytearray(f)
    g[len(g) // 2] ^= 0x10
    bad = tmp_path / "10.0.0.7_9199_nearest_neighbor_bad.jubatus"
    bad.write_bytes(bytes(g))
    with pytest.raises(nn.JubatusError, match="invalid crc32 checksum"):
        a.load("bad")
    assert a.size() == n
    with pytest.raises(nn.JubatusError, match="cannot open input file"):
        a.load("nope")


def test_light_lof_bulk_load_then_streaming_adds(oracle):
    """add_bulk = rows first, then every kdist, then every lrd; adds after it behave as usual.  The
    oracle gets the same final table state by re-scoring every row after loading (touch order
    does not matter once all kdist are fresh)."""
    cfg = json.dumps({"method": "light_lof", "converter": CONVERTER,
                      "parameter": {"nearest_neighbor_num": 3, "reverse_nearest_neighbor_num": 6, "method": "lsh",
                                    "parameter": {"hash_num": 128}}})
    n = 60
    ids, vals = w.synth_rows(w.MASTER_SEED, 900, n + 5, nnz=8, nfeat_log2=6)
    bulk = nn.Anomaly(cfg)
    got_ids = bulk.add_bulk([datum_of(ids[r], vals[r]) for r in range(n)])
    assert got_ids == [str(r) for r in range(n)]
    # reference state: a second server loaded row by row, then the same two global passes
    # (kdist of all rows, then lrd of all rows) emulated through its own parameters
    seq = nn.Anomaly(cfg)
    for r in range(n):
        seq.add(datum_of(ids[r], vals[r]))
    # the engines hold the same rows, so neighbor lists agree; kdist must equal the k-th neighbor distance
    eng = nn.NearestNeighbor(config("lsh", 128))
    for r in range(n):
        eng.set_row(str(r), datum_of(ids[r], vals[r]))
    kd = {}
    for r in range(n):
        nb = eng.neighbor_row_from_id(str(r), 3)
        kd[str(r)] = nb[-1][1]
    for r in (0, 13, 59):
        nb = eng.neighbor_row_from_id(str(r), 3)
        want_lrd = len(nb) / sum(max(d, kd[i]) for i, d in nb) if sum(max(d, kd[i]) for i, d in nb) else float("inf")
        got_kd, got_lrd = bulk.lof_params(str(r))
        assert got_kd == kd[str(r)]
        assert np.isclose(got_lrd, want_lrd, rtol=1e-5) or (np.isinf(got_lrd) and np.isinf(want_lrd))
    # streaming adds continue the id sequence and return finite-or-inf scores like the row-by-row server
    for r in range(n, n + 5):
        rid, score = bulk.add(datum_of(ids[r], vals[r]))
        assert rid == str(r) and score == score
    assert bulk.size() == n + 5


def test_recommender_and_light_lof_with_unlearner():
    """config/recommender/nearest_neighbor_recommender_euclid_lsh_unlearn_lru.json and
    config/anomaly/light_lof_unlearn_lru.json: the model never holds more than max_size rows, the
    least recently touched row goes first, and it leaves the device table as well."""
    cfg = json.dumps({"method": "nearest_neighbor_recommender", "converter": CONVERTER,
                      "parameter": {"method": "euclid_lsh", "parameter": {"hash_num": 64},
                                    "unlearner": "lru", "unlearner_parameter": {"max_size": 5}}})
    rec = nn.Recommender(cfg)
    ids, vals = w.synth_rows(w.MASTER_SEED, 300, 12, nnz=10, nfeat_log2=7)
    for r in range(5):
        rec.update_row("u%d" % r, datum_of(ids[r], vals[r]))
    rec.update_row("u0", datum_of(ids[0][:3], vals[0][:3]))          # touch u0: u1 is now the oldest
    rec.update_row("u5", datum_of(ids[5], vals[5]))                   # evicts u1
    assert sorted(rec.get_all_rows()) == ["u0", "u2", "u3", "u4", "u5"] and rec.size() == 5
    assert rec.similar_row_from_id("u1", 3) == []
    assert rec.similar_row_from_id("u5", 1)[0][0] == "u5"
    rec.clear_row("u2")                                               # frees a place: nobody is evicted next
    rec.update_row("u6", datum_of(ids[6], vals[6]))
    assert sorted(rec.get_all_rows()) == ["u0", "u3", "u4", "u5", "u6"]
    for r in range(7, 12):
        rec.update_row("u%d" % r, datum_of(ids[r], vals[r]))
    assert sorted(rec.get_all_rows()) == ["u10", "u11", "u7", "u8", "u9"]
    # the datum query only ever sees rows that are still held
    got = rec.similar_row_from_datum(datum_of(ids[3], vals[3]), 10)
    assert len(got) == 5 and all(g[0] in ("u7", "u8", "u9", "u10", "u11") for g in got)

    acfg = json.dumps({"method": "light_lof", "converter": CONVERTER,
                       "parameter": {"nearest_neighbor_num": 3, "reverse_nearest_neighbor_num": 6, "method": "lsh",
                                     "parameter": {"hash_num": 128},
                                     "unlearner": "lru", "unlearner_parameter": {"max_size": 20}}})
    an = nn.Anomaly(acfg)
    ids, vals = w.synth_rows(w.MASTER_SEED, 500, 60, nnz=8, nfeat_log2=6)
    for r in range(60):
        rid, score = an.add(datum_of(ids[r], vals[r]))
        assert rid == str(r) and score == score
        assert an.size() == min(r + 1, 20)
    assert an.lof_params("39") is None and an.lof_params("40") is not None   # rows 0..39 were unlearned
    assert sorted(an.get_all_rows(), key=int) == [str(r) for r in range(40, 60)]
    with pytest.raises(nn.JubatusError):
        nn.Anomaly(acfg.replace('"lru"', '"nope"'))
    with pytest.raises(nn.JubatusError):
        nn.Anomaly(acfg.replace('{"max_size": 20}', '{"max_size": 0}'))


def test_classifier_nn_method(oracle):
    """jubaclassifier method "NN" (config/classifier/nn.json): votes exp(-local_sensitivity * distance)
    of the nearest_neighbor_num nearest training rows, checked against the same engine queried
    directly and against the CPU oracle's engine holding the same rows (an independent path from
    datum to neighbours: its own signatures, its own ranking)."""
    import math
    k, alpha, hash_num = 8, 1.0, 128
    cfg = json.dumps({"method": "NN", "converter": CONVERTER,
                      "parameter": {"method": "euclid_lsh", "parameter": {"hash_num": hash_num},
                                    "nearest_neighbor_num": k, "local_sensitivity": alpha}})
    clf = nn.Classifier(cfg)
    eng = nn.NearestNeighbor(config("euclid_lsh", hash_num))
    ora = oracle.OracleNN("euclid_lsh", hash_num)
    ids, vals = w.synth_rows(w.MASTER_SEED, 800, 90, nnz=10, nfeat_log2=7)
    labels = ["spam", "ham", "eggs"]
    assert clf.set_label("unused") is True and clf.set_label("unused") is False
    serial = 0
    for r in range(80):
        lab = labels[r % 3]
        assert clf.train([(lab, datum_of(ids[r], vals[r]))]) == 1
        eng.set_row("%s_%d" % (lab, serial), datum_of(ids[r], vals[r]))
        ora.set_row("%s_%d" % (lab, serial), sfv_of(ids[r], vals[r]))
        serial += 1
    assert clf.get_labels() == {"spam": 27, "ham": 27, "eggs": 26, "unused": 0}
    qrows = list(range(80, 90)) + [4]
    queries = [datum_of(ids[r], vals[r]) for r in qrows]
    got = clf.classify(queries)
    assert len(got) == len(queries)
    for r, d, res in zip(qrows, queries, got):
        want = {"spam": 0.0, "ham": 0.0, "eggs": 0.0, "unused": 0.0}
        for rid, dist in eng.neighbor_row_from_datum(d, k):
            want[rid.rsplit("_", 1)[0]] += math.exp(-alpha * dist)
        assert [lab for lab, _ in res] == sorted(want)              # every known label, label order
        for lab, score in res:
            assert abs(score - want[lab]) <= 1e-5 * max(1.0, want[lab]), (lab, score, want[lab])
        # the oracle's engine: same neighbours, distances within the tolerance of north_star (1e-5 relative)
        want_o = {"spam": 0.0, "ham": 0.0, "eggs": 0.0, "unused": 0.0}
        for rid, dist in ora.neighbor_row_from_datum(sfv_of(ids[r], vals[r]), k):
            want_o[rid.rsplit("_", 1)[0]] += math.exp(-alpha * dist)
        for lab, score in res:
            assert abs(score - want_o[lab]) <= 1e-4 * max(1.0, want_o[lab]), (lab, score, want_o[lab])
    # a stored datum votes for its own label with weight exp(0) = 1 at least
    own = dict(got[-1])
    assert own[labels[4 % 3]] >= 1.0 - 1e-6
    # delete_label removes the label and its rows; clear resets everything
    assert clf.delete_label("ham") is True and clf.delete_label("ham") is False
    assert clf.get_labels() == {"spam": 27, "eggs": 26, "unused": 0}
    assert clf.size() == 53
    assert all(lab != "ham" for lab, _ in clf.classify(queries[:1])[0])
    clf.clear()
    assert clf.get_labels() == {} and clf.size() == 0
    for bad in ({"method": "AROW"}, {"parameter": {"method": "lsh", "parameter": {"hash_num": 64}, "local_sensitivity": 1}},
                {"parameter": {"method": "lsh", "parameter": {"hash_num": 64}, "nearest_neighbor_num": 0, "local_sensitivity": 1}}):
        c = json.loads(cfg)
        c.update(bad)
        with pytest.raises(nn.JubatusError):
            nn.Classifier(json.dumps(c))


def test_default_config_bigram_idf_end_to_end(tmp_path):
    """config/nearest_neighbor/default.json's converter (bigram tokens, tf x idf) in front of the
    engine: stored rows move the document counts, the counts travel with the model file."""
    conv = {"string_filter_types": {}, "string_filter_rules": [], "num_filter_types": {}, "num_filter_rules": [],
            "string_types": {"bigram": {"method": "ngram", "char_num": "2"}},
            "string_rules": [{"key": "*", "type": "bigram", "sample_weight": "tf", "global_weight": "idf"}],
            "num_types": {}, "num_rules": [{"key": "*", "type": "num"}]}
    cfg = json.dumps({"method": "lsh", "parameter": {"hash_num": 256}, "converter": conv})
    a = nn.NearestNeighbor(cfg)
    a.set_storage(str(tmp_path), "127.0.0.1", 9199)
    texts = ["the quick brown fox", "the quick brown dog", "lorem ipsum dolor sit amet", "jubatus online learning",
             "nearest neighbor search", "the lazy dog sleeps", "online machine learning framework", "quick sort"]
    # a row is hashed with the weights of the moment it is stored (the first one sees an empty
    # model: idf 0, all-zero signature), so every row is stored a second time once the counts exist
    for _ in range(2):
        for i, t in enumerate(texts):
            a.set_row("t%d" % i, nn.Datum().add_string("text", t))
    got = a.neighbor_row_from_datum(nn.Datum().add_string("text", "the quick brown fox jumps"), 3)
    assert got[0][0] == "t0" and got[1][0] == "t1"
    sim = a.similar_row_from_datum(nn.Datum().add_string("text", "online learning"), 2)
    assert {sim[0][0], sim[1][0]} == {"t3", "t6"}
    before = a.neighbor_row_from_datum(nn.Datum().add_string("text", "machine learning"), 4)
    a.save("w")
    b = nn.NearestNeighbor(cfg)
    b.set_storage(str(tmp_path), "127.0.0.1", 9199)
    b.load("w")
    assert b.neighbor_row_from_datum(nn.Datum().add_string("text", "machine learning"), 4) == before   # same idf after load
    b.clear()
    assert b.size() == 0


@pytest.mark.parametrize("method,hash_num,seed", [("lsh", 128, 1), ("euclid_lsh", 64, 2), ("minhash", 64, 3), ("lsh", 512, 4)])
def test_random_sequence_of_rpcs_matches_the_oracle_engine(oracle, method, hash_num, seed):
    """A few hundred randomly interleaved set_row (new and overwriting), delete_row, *_from_id and
    *_from_datum calls: ids, order and the row set must track the oracle's column_table
    (swap-delete moves the last row into the hole, which changes tie order afterwards)."""
    rng = np.random.RandomState(seed)
    srv = nn.NearestNeighbor(config(method, hash_num))
    ora = oracle.OracleNN(method, hash_num)
    pool_ids, pool_vals = w.synth_rows(w.MASTER_SEED, 5000 + seed, 400, nnz=6, nfeat_log2=5)   # few features: many ties
    live = []
    for step in range(350):
        op = rng.randint(0, 10)
        r = int(rng.randint(0, 400))
        if op <= 4 or not live:                       # set_row: new id or overwrite
            rid = "r%d" % int(rng.randint(0, 120))
            srv.set_row(rid, datum_of(pool_ids[r], pool_vals[r]))
            ora.set_row(rid, sfv_of(pool_ids[r], pool_vals[r]))
            if rid not in live:
                live.append(rid)
        elif op == 5:                                  # delete_row
            rid = live[int(rng.randint(0, len(live)))]
            assert srv.delete_row(rid) is True
            ora.delete_row(rid)
            live.remove(rid)
            assert srv.delete_row(rid) is False      # already gone
        elif op <= 7:                                  # query by id
            rid = live[int(rng.randint(0, len(live)))]
            k = int(rng.randint(1, 12))
            if op == 6:
                same(srv.neighbor_row_from_id(rid, k), ora.neighbor_row_from_id(rid, k), rel=1e-5)
            else:
                same(srv.similar_row_from_id(rid, k), ora.similar_row_from_id(rid, k), rel=1e-5)
        else:                                          # query by datum
            k = int(rng.randint(1, 12))
            d, s = datum_of(pool_ids[r], pool_vals[r]), sfv_of(pool_ids[r], pool_vals[r])
            if op == 8:
                same(srv.neighbor_row_from_datum(d, k), ora.neighbor_row_from_datum(s, k), rel=1e-5)
            else:
                same(srv.similar_row_from_datum(d, k), ora.similar_row_from_datum(s, k), rel=1e-5)
        if step % 50 == 49:
            assert srv.get_all_rows() == ora.get_all_rows() and srv.size() == len(live)
    assert srv.neighbor_row_from_id("never-stored", 3) == []


def test_mix_carries_the_converter_weights_too():
    """Two replicas store different texts, MIX, and then weigh a query's features alike: the idf
    counts are a mixable next to the table (every replica's diff summed into every master)."""
    conv = {"string_types": {"bigram": {"method": "ngram", "char_num": "2"}},
            "string_rules": [{"key": "*", "type": "bigram", "sample_weight": "tf", "global_weight": "idf"}],
            "num_types": {}, "num_rules": []}
    cfg = json.dumps({"method": "lsh", "parameter": {"hash_num": 128}, "converter": conv})
    a, b = nn.NearestNeighbor(cfg, my_id="A"), nn.NearestNeighbor(cfg, my_id="B")
    ta = ["alpha beta", "beta gamma", "gamma delta", "alpha delta"]
    tb = ["omega psi", "psi chi", "chi phi", "alpha omega", "beta phi"]
    for i, t in enumerate(ta):
        a.set_row("a%d" % i, nn.Datum().add_string("t", t))
    for i, t in enumerate(tb):
        b.set_row("b%d" % i, nn.Datum().add_string("t", t))
    a.mix_with(b)
    assert sorted(a.get_all_rows()) == sorted(b.get_all_rows()) and a.size() == 9
    # same rows (signatures travel as stored) and same document counts -> same scores on both
    # (row order differs between the replicas, so ties may come out in another order)
    for text in ("alpha beta gamma", "psi phi", "delta omega"):
        q = nn.Datum().add_string("t", text)
        ra, rb = a.similar_row_from_datum(q, 9), b.similar_row_from_datum(q, 9)
        assert sorted(ra, key=lambda x: (-x[1], x[0])) == sorted(rb, key=lambda x: (-x[1], x[0]))
        assert [s for _, s in ra] == [s for _, s in rb]
    # a second MIX with nothing new changes nothing
    before = a.similar_row_from_datum(nn.Datum().add_string("t", "alpha"), 9)
    a.mix_with(b)
    assert a.similar_row_from_datum(nn.Datum().add_string("t", "alpha"), 9) == before
