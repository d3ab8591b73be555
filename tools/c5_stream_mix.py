"""BASELINE config 5 in miniature, on real GPUs over NCCL (run under torchrun / gpurun --gpus N):
every rank is a replica (jubaanomaly-style: the whole table on each), rows stream in on the rank
that owns them, and every `--interval` adds the replicas MIX by all-gathering their versioned row
diffs as device records (jubatus_b200.dist.ReplicaMixer).  At the end all replicas must hold the
same rows and answer a query batch identically, and equal to the CPU oracle over the union.

    python -m torch.distributed.run --nproc-per-node 2 --master-addr 127.0.0.1 tools/c5_stream_mix.py
"""
import argparse
import json
import os
import sys
import time

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
from jubatus_b200 import Engine, workload, dist as jd  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--rows", type=int, default=20000)
    ap.add_argument("--hash-num", type=int, default=256)
    ap.add_argument("--interval", type=int, default=512)
    ap.add_argument("--k", type=int, default=10)
    a = ap.parse_args()
    rank, world = int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    s32, k64 = workload.feature_tables(20)
    eng = Engine("lsh", a.hash_num, local)
    table = jd.ReplicaTable(jd.EngineColumns(eng), my_id="server%d" % rank)
    mixer = jd.ReplicaMixer(table)
    ids, vals = workload.synth_rows(workload.MASTER_SEED, 0, a.rows, nnz=100)
    indptr, seed32, _, v = workload.to_csr(ids, vals, s32, k64)
    sigs, _ = eng.hash_csr(indptr, seed32, None, v)          # signatures of every datum (device kernel)
    t0 = time.time()
    mixes = 0
    batch = a.interval
    for r0 in range(0, a.rows, batch * world):
        mine = [g for g in range(r0 + rank * batch, min(a.rows, r0 + (rank + 1) * batch))]
        if mine:
            table.add(["row%d" % g for g in mine], sigs[mine])
        mixer.mix()                                          # linear_mixer: gather diffs, apply everywhere
        mixes += 1
    torch.cuda.synchronize()
    secs = time.time() - t0
    # every replica now holds every row
    assert table.size() == a.rows, (rank, table.size())
    order = [table.index["row%d" % g] for g in range(a.rows)]
    got_sigs, _ = eng.get_rows_sig(np.array(order, dtype=np.uint64))
    assert np.array_equal(got_sigs, sigs), "replica content differs from the union"
    q = sigs[:32]
    idx, score, count = eng.query_sig(q, None, a.k)
    names = np.array(table.keys)[idx.astype(np.int64)]
    gathered = [None] * world
    dist.all_gather_object(gathered, (names.tolist(), score.tolist()))
    if rank == 0:
        # replicas store rows in different orders; ids and distances must still agree wherever no tie decides
        import oracle_lib
        oidx, oscore, _, _ = oracle_lib.scan_batch("lsh", q, None, sigs, None, a.hash_num, a.k, threads=4)
        for g in gathered:
            assert np.array_equal(np.array(g[1], dtype=np.float32), oscore), "distances differ from the oracle"
        print(json.dumps({"replicas": world, "rows": a.rows, "hash_num": a.hash_num, "mix_rounds": mixes,
                          "adds_per_s_incl_mix": round(a.rows / secs), "record_bytes": eng.diff_record_bytes(),
                          "check": "all replicas hold identical rows; query distances equal the CPU oracle"}))
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
