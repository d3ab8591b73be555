"""N > 1 plumbing on CPU: world_size-2 gloo process groups exercise the sharding arithmetic,
the candidate all-gather layout and the MIX exchange + version arbitration of
jubatus_b200/dist.py.  The per-rank device work is stood in for by the oracle (shard-local
ranking) and a numpy column store; the GPU versions of the same steps are covered by
tests/test_sharded_gpu.py."""
import os
import socket
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _init(rank, world, port):
    sys.path.insert(0, ROOT)
    sys.path.insert(0, HERE)
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)


def merge_parts_numpy(all_k, all_i, k):
    """(key, global index) ascending over all parts, sentinels dropped -- what merge_parts_kernel does."""
    parts, nq, _ = all_k.shape
    out = []
    for q in range(nq):
        cand = [(int(all_k[p, q, j]) & 0xFFFFFFFF, int(all_i[p, q, j]) & 0xFFFFFFFFFFFFFFFF)
                for p in range(parts) for j in range(all_k.shape[2])]
        cand = sorted(c for c in cand if c[1] != 0xFFFFFFFFFFFFFFFF)[:k]
        out.append(cand)
    return out


def _sharded_worker(rank, world, port, q):
    _init(rank, world, port)
    import oracle_lib
    from jubatus_b200 import dist as jd
    rng = np.random.RandomState(0)                       # same data on every rank
    hash_num, n, nq, k = 128, 3001, 5, 20
    rows = rng.randint(0, 2 ** 62, size=(n, 2), dtype=np.int64).astype(np.uint64)
    rows[100:900] = rows[5]                              # ties that straddle the shards
    queries = np.concatenate([rows[5:6], rng.randint(0, 2 ** 62, size=(nq - 1, 2), dtype=np.int64).astype(np.uint64)])
    mine = np.arange(rank, n, world)
    assert len(mine) == jd.local_count(n, rank, world)
    assert all(jd.shard_of(g, world) == rank and jd.global_index(jd.local_slot(g, world), rank, world) == g
               for g in mine[:50])
    lidx, lscore, lcount, _ = oracle_lib.scan_batch("lsh", queries, None, rows[mine], None, hash_num, k)
    ck = torch.full((nq, k), -1, dtype=torch.int32)      # 0xFFFFFFFF
    ci = torch.full((nq, k), -1, dtype=torch.int64)      # UINT64_MAX
    for qi in range(nq):
        c = int(lcount[qi])
        ck[qi, :c] = torch.from_numpy(np.rint(lscore[qi, :c] * hash_num).astype(np.int32))
        ci[qi, :c] = torch.from_numpy(jd.global_index(lidx[qi, :c].astype(np.int64), rank, world))
    all_k, all_i = jd.gather_candidates(ck, ci)
    assert tuple(all_k.shape) == (world, nq, k)
    merged = merge_parts_numpy(all_k.numpy(), all_i.numpy(), k)
    widx, wscore, wcount, _ = oracle_lib.scan_batch("lsh", queries, None, rows, None, hash_num, k)
    ok = all([m[1] for m in merged[qi]] == list(widx[qi]) and
             [m[0] for m in merged[qi]] == list(np.rint(wscore[qi] * hash_num).astype(int)) for qi in range(nq))
    if rank == 0:
        q.put(bool(ok))
    dist.barrier()
    dist.destroy_process_group()


class NumpyColumns(object):
    """CPU stand-in for EngineColumns: same record format as diff_pack/apply_kernel."""

    def __init__(self, w64):
        self.w64 = w64
        self.device = torch.device("cpu")
        self.record_bytes = 8 + 8 * w64 + 8
        self.sigs = np.zeros((0, w64), dtype=np.uint64)
        self.norms = np.zeros(0, dtype=np.float32)

    def reserve(self, rows):
        if rows > len(self.sigs):
            extra = rows - len(self.sigs)
            self.sigs = np.concatenate([self.sigs, np.zeros((extra, self.w64), dtype=np.uint64)])
            self.norms = np.concatenate([self.norms, np.zeros(extra, dtype=np.float32)])

    def set_rows_sig(self, rows, sigs, norms=None):
        self.reserve(int(max(rows)) + 1)
        for i, r in enumerate(rows):
            self.sigs[int(r)] = sigs[i]
            self.norms[int(r)] = 0.0 if norms is None else norms[i]

    def diff_pack(self, rows):
        rec = np.zeros((len(rows), self.record_bytes), dtype=np.uint8)
        for i, r in enumerate(rows):
            rec[i, :8] = np.array([r], dtype=np.uint64).view(np.uint8)
            rec[i, 8:8 + 8 * self.w64] = self.sigs[int(r)].view(np.uint8)
            rec[i, 8 + 8 * self.w64:12 + 8 * self.w64] = np.array([self.norms[int(r)]], dtype=np.float32).view(np.uint8)
        return torch.from_numpy(rec)

    def diff_apply(self, records):
        rec = records.numpy()
        for i in range(rec.shape[0]):
            r = int(rec[i, :8].copy().view(np.uint64)[0])
            self.sigs[r] = rec[i, 8:8 + 8 * self.w64].copy().view(np.uint64)
            self.norms[r] = rec[i, 8 + 8 * self.w64:12 + 8 * self.w64].copy().view(np.float32)[0]


def _mix_worker(rank, world, port, q):
    _init(rank, world, port)
    from jubatus_b200 import dist as jd
    w64 = 2
    t = jd.ReplicaTable(NumpyColumns(w64), my_id="server%d" % rank)
    mixer = jd.ReplicaMixer(t)

    def sig(tag):
        return np.array([[tag, tag * 7 + 1]], dtype=np.uint64)

    # round 1: disjoint rows on each replica + one key written on both
    for j in range(3):
        t.add(["r%d_%d" % (rank, j)], sig(100 * rank + j))
    t.add(["shared"], sig(1000 + rank))
    applied = mixer.mix()
    state1 = {k: (tuple(t.engine.sigs[t.index[k]]), t.versions[t.index[k]]) for k in t.keys}
    # round 2: nothing changed -> nothing travels
    nothing = mixer.mix()
    # round 3: rank 1 updates a row that rank 0 created; its clock is now past everything it has seen
    if rank == 1:
        t.add(["r0_0"], sig(4242))
    mixer.mix()
    state3 = {k: (tuple(t.engine.sigs[t.index[k]]), t.versions[t.index[k]]) for k in t.keys}
    states = [None] * world
    dist.all_gather_object(states, (state1, state3, applied, nothing, t.clock))
    if rank == 0:
        q.put(states)
    dist.barrier()
    dist.destroy_process_group()


def _run(worker):
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    out = q.get(timeout=120)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    return out


@pytest.mark.timeout(180)
def test_sharded_candidates_allgather_and_merge_equal_single_table():
    assert _run(_sharded_worker) is True


@pytest.mark.timeout(180)
def test_mix_allgather_of_versioned_row_diffs():
    states = _run(_mix_worker)
    (a1, a3, a_applied, a_nothing, a_clock), (b1, b3, b_applied, b_nothing, b_clock) = states
    # after one round both replicas hold the same 7 keys with the same contents and versions
    assert set(a1) == set(b1) and len(a1) == 7
    assert a1 == b1
    # "shared": both wrote it at local clock 3; equal clocks -> the incoming row wins on each side in rank order,
    # so both end with the last one applied (rank 1's)
    assert a1["shared"][0] == (1001, 1001 * 7 + 1)
    assert a_nothing == 0 and b_nothing == 0
    # the update made on replica 1 reached replica 0 (newer clock wins) and nothing else changed
    assert a3["r0_0"][0] == (4242, 4242 * 7 + 1) and a3 == b3
    assert a3["r0_0"][1][0] == "server1"
    assert all(a3[k] == a1[k] for k in a1 if k != "r0_0")
    assert a_clock == b_clock
