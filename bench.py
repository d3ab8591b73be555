#!/usr/bin/env python
"""Headline benchmark: neighbor queries/sec @ 100M x 512-bit LSH rows, k=100 (BASELINE.json).

    python bench.py [--gpus N] [--steps K] [--warmup W]               # this repo's CUDA path
    python bench.py --impl reference [--gpus N] --steps K --warmup W  # CPU engine (oracle) on host cores

A step = one batch of --nq neighbor queries (k=100) against the full table.  The table (10^8 rows x
512 bits = 6.4 GB; 50x the 126 MB L2, so nothing is cache-resident between passes) is built from the
synthetic datums of SURVEY.md 8(d) -- splitmix64, seed 0x4A554241, 100 distinct ascending feature ids
of 2^20 per row -- hashed on the device by the signature kernel, and is row-sharded round-robin over
the N GPUs (each rank generates and hashes only the rows it keeps).  Queries come from the disjoint
seed stream with 1 % exact and 1 % near duplicates of stored rows.

`value`  device-resident path: jubann_query_sig_dev (CUDA events on the launching stream, max over
         ranks); on N > 1 GPUs that one call is collective inside libjubann.so (shard-local scan, one
         packed NCCL all-gather, merge).
`e2e`    the host-buffer call an adapter makes, jubann_query_csr: pinned host CSR of the query datums
         in, ids / scores / counts out, copies and hashing inside the timed region; collective on N > 1.
`parity` in the same run: the table the GPU holds is brought to the host and the CPU engine (oracle)
         ranks the same queries against it; ids, order and score bit patterns must be identical, and a
         sample of stored signatures is re-derived from the datums by the oracle.  The run fails
         otherwise.  That CPU scan is also the `cpu_baseline` sample.
`configs` short runs of BASELINE.json's other configurations (c2..c5), each with its own parity check
         and both roofline fractions (HBM and POPC issue).
Prints ONE JSON line on rank 0.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "neighbor queries/sec @100Mx512-bit LSH rows, k=100"
UNIT = "queries/s"
NNZ = 100
NFEAT_LOG2 = 20


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    # workload knobs (defaults = the headline configuration)
    ap.add_argument("--method", default="lsh")
    ap.add_argument("--rows", type=int, default=100_000_000)
    ap.add_argument("--hash-num", type=int, default=512)
    ap.add_argument("--k", type=int, default=100)
    ap.add_argument("--nq", type=int, default=64, help="queries per step (batch)")
    ap.add_argument("--qp", type=int, default=2,
                    help="queries sharing one streaming pass in the headline run (2 = HBM-bound regime; 0 = library default)")
    ap.add_argument("--popc-variant", type=int, default=-1)
    ap.add_argument("--scan-variant", type=int, default=0)
    ap.add_argument("--configs", default="c2,c3,c4,c5", help="comma-separated subset of c2,c3,c4,c5 ('' = none)")
    ap.add_argument("--config-scale", type=float, default=1.0, help="shrink the rows of c2..c5 (development runs)")
    ap.add_argument("--opt", action="append", default=[], help="library option name=value for the headline engine (tuning runs)")
    ap.add_argument("--skip-cpu", action="store_true", help="omit the cpu_baseline / parity leg")
    ap.add_argument("--parity-queries", type=int, default=16)
    return ap.parse_args()


def workload_name(method, hash_num, rows, nq, k):
    return "%s hash_num=%d, %d rows, %d-query batches, k=%d" % (method, hash_num, rows, nq, k)


# --------------------------------------------------------------------------------------
# clocks
# --------------------------------------------------------------------------------------
class ClockSampler(object):
    FIELDS = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,"
              "clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
              "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index):
        self.gpu_index = gpu_index
        self.lines = []
        self.proc = None

    def start(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", "-i", str(self.gpu_index), "--query-gpu=" + self.FIELDS, "--format=csv,noheader,nounits",
                 "-lms", "100"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._pump, daemon=True)
            self.thread.start()
        except Exception:
            self.proc = None

    def _pump(self):
        for line in self.proc.stdout:
            self.lines.append((time.time(), line.strip()))

    def stop(self, t0, t1):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        sm, mx, reasons = [], [], set()
        for ts, line in self.lines:
            if ts < t0 - 0.05 or ts > t1 + 0.05:
                continue
            f = [x.strip() for x in line.split(",")]
            if len(f) < 9:
                continue
            try:
                sm.append(float(f[1])); mx.append(float(f[2]))
            except ValueError:
                continue
            for name, val in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), f[5:9]):
                if val.lower().startswith("active"):
                    reasons.add(name)
        if not sm:  # region shorter than the sampling period: take the nearest samples
            for ts, line in self.lines[-3:]:
                f = [x.strip() for x in line.split(",")]
                try:
                    sm.append(float(f[1])); mx.append(float(f[2]))
                except (ValueError, IndexError):
                    pass
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "samples": len(sm), "reasons": sorted(reasons)}


# --------------------------------------------------------------------------------------
# CPU legs (oracle = the reference algorithm restated; test infrastructure, timed here only)
# --------------------------------------------------------------------------------------
def oracle():
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import oracle_lib
    return oracle_lib


def cpu_rows_that_fit(rows, w64):
    try:
        avail = int([l for l in open("/proc/meminfo") if l.startswith("MemAvailable")][0].split()[1]) * 1024
    except Exception:
        avail = 8 << 30
    limit = int(avail * 0.5) // (w64 * 8)
    return min(rows, max(1_000_000, limit))


def run_reference(a):
    """--impl reference: the CPU engine (oracle port; the reference itself cannot be built here,
    DESIGN.md) with every host thread, on the same workload shape; each step is a bounded sample.
    The table is uniform random signature words (LSH bits of the synthetic datums are uniform too;
    hashing 10^8 datums on the CPU would take longer than the run may) -- the arm that proves
    agreement on identical data is the CUDA arm's `parity` leg."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    from jubatus_b200 import workload
    ol = oracle()
    threads = os.cpu_count() or 1
    w64 = (a.hash_num + 63) // 64
    n_cpu = cpu_rows_that_fit(a.rows, w64)
    table = np.empty((n_cpu, w64), dtype=np.uint64)
    ol.fill_random_words(workload.MASTER_SEED, n_cpu * w64, threads=threads, out=table.reshape(-1))
    euclid = a.method == "euclid_lsh"
    norms = np.ones(n_cpu, dtype=np.float32) if euclid else None

    def step(nq, seed):
        q = workload.random_signatures(seed, nq, w64)
        q[0] = table[(seed * 7919) % n_cpu]
        qn = np.ones(nq, dtype=np.float32) if euclid else None
        return ol.scan_batch(a.method, q, qn, table, norms, a.hash_num, a.k, threads=threads, flavour="v3")[3]

    t1 = step(1, 1)
    budget = 150.0
    nq_step = int(max(1, min(a.nq, budget / max(1, a.steps + a.warmup) / max(t1, 1e-4))))
    for w in range(a.warmup):
        step(nq_step, 100 + w)
    total = 0.0
    for s in range(a.steps):
        total += step(nq_step, 200 + s)
    scale = n_cpu / float(a.rows)
    qps = a.steps * nq_step / total * scale
    sample = "each step = %d of the batch's %d queries over %d rows (k=%d), threads=%d, build O3-x86-64-v3" % (
        nq_step, a.nq, n_cpu, a.k, threads)
    if n_cpu != a.rows:
        sample += "; table shrunk to fit host RAM, value scaled by %.3f" % scale
    line = {
        "impl": "reference", "metric": METRIC, "value": qps, "unit": UNIT, "n_gpus": a.gpus, "steps": a.steps,
        "warmup": a.warmup, "ms_per_step": total / a.steps * 1e3, "higher_is_better": True, "scaling": "strong",
        "vs_baseline": None, "dtype": "u64", "data": "synthetic (uniform random signature words)",
        "config": {"workload": workload_name(a.method, a.hash_num, a.rows, a.nq, a.k),
                   "note": "CPU engine = oracle port of jubatus_core 0.3.2 (reference not buildable here)"},
        "cpu_baseline": {"value": qps, "unit": UNIT, "cores": threads, "kind": "port", "sample": sample},
        "e2e": {"value": qps, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }
    emit(line)


# --------------------------------------------------------------------------------------
# CUDA arm
# --------------------------------------------------------------------------------------
class Env(object):
    """Process-group plumbing: torch.distributed carries the barrier, the max over ranks and the
    128-byte NCCL id of the library's own communicator; no table or query data goes through it."""

    def __init__(self, a):
        import torch
        import torch.distributed as dist
        self.torch, self.dist = torch, dist
        self.world = int(os.environ.get("WORLD_SIZE", "1"))
        self.rank = int(os.environ.get("RANK", "0"))
        self.local_rank = int(os.environ.get("LOCAL_RANK", "0"))
        if self.world > 1:
            a.gpus = self.world
        if not torch.cuda.is_available():
            raise RuntimeError("bench.py needs a CUDA device (no CPU fallback)")
        torch.cuda.set_device(self.local_rank)
        self.dev = torch.device("cuda", self.local_rank)
        if self.world > 1:
            dist.init_process_group("nccl", device_id=self.dev)
        self.stream = torch.cuda.current_stream().cuda_stream
        self.peaks = {}
        try:
            self.peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        except Exception:
            pass
        self.hbm_peak = float(self.peaks.get("hbm_gbs", 6650.0))
        self.micro = {}
        try:
            self.micro = json.load(open(os.path.join(ROOT, "profiles", "microbench.json")))
        except Exception:
            pass
        from jubatus_b200 import workload
        s32, k64 = workload.feature_tables(NFEAT_LOG2)
        self.s32, self.k64 = s32, k64
        self.d_s32 = torch.from_numpy(s32.view(np.int32)).to(self.dev)
        self.d_k64 = torch.from_numpy(k64.view(np.int64)).to(self.dev)

    def barrier(self):
        self.torch.cuda.synchronize()
        if self.world > 1:
            self.dist.barrier()
            self.torch.cuda.synchronize()

    def max_over_ranks(self, x):
        if self.world == 1:
            return x
        t = self.torch.tensor([x], dtype=self.torch.float64, device=self.dev)
        self.dist.all_reduce(t, op=self.dist.ReduceOp.MAX)
        return float(t.item())


def issued_popc_per_pair(hash_num, popc_variant=3):
    """POPC instructions the scan kernels issue per (row, query) pair after the carry-save adders."""
    w32 = 2
    while w32 * 32 < hash_num:
        w32 *= 2
    if popc_variant == 0 or w32 == 2:
        return w32
    if w32 == 4:
        return 3
    if w32 == 8:
        return 4
    per16 = {1: 11, 3: 9, 4: 7}.get(popc_variant, 5)
    return per16 * (w32 // 16)


def scan_kernel_name(hash_num, qp, local_rows, num_sms=148):
    """The automatic choice of plan_scan (csrc/scan.cu): the TMA ring while a pass is memory-bound."""
    w32 = 2
    while w32 * 32 < hash_num:
        w32 *= 2
    tma = qp <= (3 if w32 >= 16 else 4) and (local_rows + 31) // 32 >= num_sms * 32
    return "scan_tma_kernel" if tma else "scan_ldg_kernel"


def build_table(env, eng, method, rows, seed, chunk=1_000_000):
    """Rows g = rank, rank + G, ... of the synthetic table: generated on the device, hashed by the
    signature kernel and stored (set_row end to end); no rank touches a row it does not keep."""
    torch = env.torch
    from jubatus_b200 import workload
    G, rank = env.world, env.rank
    minhash = method == "minhash"
    local = (rows - rank + G - 1) // G if rows > rank else 0
    indptr = torch.arange(chunk + 1, device=env.dev, dtype=torch.int64) * NNZ
    t0 = time.time()
    for j0 in range(0, local, chunk):
        n = min(chunk, local - j0)
        g = (torch.arange(j0, j0 + n, device=env.dev, dtype=torch.int64) * G + rank)
        ids, vals = workload.synth_rows_torch(seed, g, NNZ, NFEAT_LOG2)
        flat = ids.reshape(-1)
        seed32 = None if minhash else env.d_s32[flat].contiguous()
        keyh = env.d_k64[flat].contiguous() if minhash else None
        v = vals.reshape(-1).contiguous()
        eng.set_rows_csr_strided_dev(n, rank + G * j0, G, n * NNZ, indptr.data_ptr(),
                                     seed32.data_ptr() if seed32 is not None else 0,
                                     keyh.data_ptr() if keyh is not None else 0, v.data_ptr(), env.stream)
        torch.cuda.synchronize()
        del g, ids, vals, flat, seed32, keyh, v
    torch.cuda.synchronize()
    return time.time() - t0


class QuerySets(object):
    """`nsets` batches of `nq` query datums (disjoint seed stream, 1 % exact + 1 % near
    duplicates of stored rows): pinned host CSR (what the adapter hands to jubann_query_csr), a device
    copy, and their signatures resident in HBM for the device-timed region."""

    def __init__(self, env, eng, method, hash_num, rows, nq, nsets, seed):
        torch = env.torch
        from jubatus_b200 import workload
        self.nq, self.nsets = nq, nsets
        self.minhash = method == "minhash"
        self.euclid = method == "euclid_lsh"
        w64 = (hash_num + 63) // 64
        self.h, self.d_q, self.d_qn, self.exact_of = [], [], [], []
        for i in range(nsets):
            ids, vals, exact_of, _ = workload.make_queries_lazy(seed, nq, rows, NNZ, NFEAT_LOG2, query_row0=i * nq)
            indptr, seed32, keyhash, v = workload.to_csr(ids, vals, env.s32, env.k64)
            c = {"indptr": torch.from_numpy(indptr.view(np.int64)).pin_memory(),
                 "seed": torch.from_numpy(seed32.view(np.int32)).pin_memory(),
                 "key": torch.from_numpy(keyhash.view(np.int64)).pin_memory(),
                 "vals": torch.from_numpy(v).pin_memory()}
            self.h.append(c)
            dc = {k_: x.to(env.dev) for k_, x in c.items()}
            q = torch.zeros((nq, w64), dtype=torch.int64, device=env.dev)
            qn = torch.zeros(nq, dtype=torch.float32, device=env.dev)
            eng.hash_csr_dev(nq, nq * NNZ, dc["indptr"].data_ptr(), 0 if self.minhash else dc["seed"].data_ptr(),
                             dc["key"].data_ptr() if self.minhash else 0, dc["vals"].data_ptr(), q.data_ptr(), qn.data_ptr(),
                             env.stream)
            self.d_q.append(q); self.d_qn.append(qn); self.exact_of.append(exact_of)
        torch.cuda.synchronize()
        self.h2d_bytes = nq * NNZ * (4 + (8 if self.minhash else 4)) + (nq + 1) * 8


def fetch_table(env, eng, method, rows, w64):
    """The table the GPUs hold, on rank 0's host: [rows][w64] words (+ norms).  Shards of other ranks
    travel through /dev/shm files (host-side test plumbing, after the timed regions)."""
    G, rank = env.world, env.rank
    euclid = method == "euclid_lsh"
    local = (rows - rank + G - 1) // G if rows > rank else 0
    sig = np.empty((local, w64), dtype=np.uint64)
    nrm = np.empty(local, dtype=np.float32) if euclid else None
    step = 4_000_000
    for j0 in range(0, local, step):
        n = min(step, local - j0)
        idx = (np.arange(j0, j0 + n, dtype=np.uint64) * np.uint64(G) + np.uint64(rank))
        s, m = eng.get_rows_sig(idx)
        sig[j0:j0 + n] = s
        if euclid:
            nrm[j0:j0 + n] = m
    if G == 1:
        return sig, nrm
    # where the shards meet: /dev/shm if it has the room, else the temp directory; rank 0 decides for all
    choice = [None]
    if rank == 0:
        import shutil
        import tempfile
        need = int(rows * (w64 * 8 + 4) * 1.2)
        for d in ("/dev/shm", tempfile.gettempdir()):
            try:
                if shutil.disk_usage(d).free >= need:
                    choice[0] = d
                    break
            except OSError:
                pass
    env.dist.broadcast_object_list(choice, src=0)
    if choice[0] is None:
        return None, "no directory with %d bytes free to bring the shards together" % int(rows * (w64 * 8 + 4) * 1.2)
    tag = os.path.join(choice[0], "jubann_bench_%d_" % os.getppid())
    np.save(tag + "sig_%d.npy" % rank, sig)
    if euclid:
        np.save(tag + "nrm_%d.npy" % rank, nrm)
    del sig, nrm
    env.barrier()
    table = norms = None
    if rank == 0:
        table = np.empty((rows, w64), dtype=np.uint64)
        norms = np.empty(rows, dtype=np.float32) if euclid else None
        for r in range(G):
            table[r::G] = np.load(tag + "sig_%d.npy" % r, mmap_mode="r")
            if euclid:
                norms[r::G] = np.load(tag + "nrm_%d.npy" % r, mmap_mode="r")
    env.barrier()
    for suffix in ("sig_%d.npy" % rank, "nrm_%d.npy" % rank):
        try:
            os.remove(tag + suffix)
        except OSError:
            pass
    return table, norms


def check_parity(env, eng, method, hash_num, rows, k, flags, qs, got, nparity, seed, threads):
    """Rank 0: the oracle ranks the first `nparity` queries of set 0 against the GPU's own table and a
    sample of stored signatures is re-derived from the datums.  Returns (parity dict, oracle seconds)."""
    from jubatus_b200 import workload
    w64 = (hash_num + 63) // 64
    euclid = method == "euclid_lsh"
    table, norms = fetch_table(env, eng, method, rows, w64)
    if table is None and isinstance(norms, str):   # (N > 1 only) the host could not hold the exchange files: say so, do not guess
        return ({"exact": None, "skipped": norms} if env.rank == 0 else None), None, 0
    if env.rank != 0:
        return None, None, 0
    ol = oracle()
    # a few of each kind: exact duplicates, near duplicates, plain queries
    ndup = max(1, qs.nq // 100)
    sel = sorted(set([0, 1, ndup, ndup + 1] + list(range(2 * ndup, 2 * ndup + nparity))))
    sel = np.array([i for i in sel if i < qs.nq][:nparity], dtype=np.int64)
    nparity = len(sel)
    q = qs.d_q[0].cpu().numpy().view(np.uint64)[sel]
    qn = qs.d_qn[0].cpu().numpy()[sel] if euclid else None
    widx, wscore, wcount, secs = ol.scan_batch(method, q, qn, table, norms, hash_num, k, threads=threads,
                                               similar=bool(flags & 1), flavour="v3")
    gidx, gscore, gcount = got[0][sel], got[1][sel], got[2][sel]
    exact_of = qs.exact_of[0][sel]
    exact = True
    for i in range(nparity):
        c = int(wcount[i])
        exact = exact and int(gcount[i]) == c and np.array_equal(gidx[i, :c], widx[i, :c]) and \
            np.array_equal(gscore[i, :c].view(np.uint32), wscore[i, :c].view(np.uint32))
    # planted exact duplicates come back first at distance 0 (score 1 / -0 for similarity flags)
    planted = [int(r) for r in exact_of if r >= 0]
    found = all(int(r) in gidx[i, :max(1, int(gcount[i]))] for i, r in enumerate(exact_of) if r >= 0)
    # signatures: the oracle re-derives a sample of stored rows from their datums
    sample = np.unique(np.concatenate([np.arange(0, min(rows, 256)), (np.arange(256) * 2654435761) % rows]).astype(np.int64))
    ids = np.zeros((len(sample), NNZ), dtype=np.uint32)
    vals = np.zeros((len(sample), NNZ), dtype=np.float32)
    for j, r in enumerate(sample):
        i1, v1 = workload.synth_rows(seed, int(r), 1, NNZ, NFEAT_LOG2)
        ids[j], vals[j] = i1[0], v1[0]
    indptr, seed32, keyhash, v = workload.to_csr(ids, vals, env.s32, env.k64)
    want, wnorm = ol.signatures_csr(method, hash_num, indptr, None if method == "minhash" else seed32,
                                    keyhash if method == "minhash" else None, v, threads=threads)
    diff = np.unpackbits((want ^ table[sample]).view(np.uint8)).sum()
    agreement = 1.0 - float(diff) / (len(sample) * hash_num)
    norms_equal = bool(np.array_equal(wnorm.view(np.uint32), norms[sample].view(np.uint32))) if euclid else None
    parity = {"queries": int(nparity), "rows": int(rows), "k": int(k), "exact": bool(exact and found),
              "compared": "ids, order and score bit patterns against the CPU engine (oracle) on the table the GPUs hold",
              "planted_duplicates_found": len(planted) if found else 0,
              "signature_bit_agreement": agreement, "signature_rows_checked": int(len(sample))}
    if euclid:
        parity["norms_bit_exact"] = norms_equal
    if not parity["exact"] or agreement < 0.9999 or norms_equal is False:
        raise RuntimeError("PARITY FAILURE: %s" % json.dumps(parity))
    del table, norms
    return parity, secs, nparity


def time_device_steps(env, eng, qs, k, flags, steps, warmup, outs, sampler=None):
    torch = env.torch
    d_idx, d_score, d_count = outs

    def step(i):
        s = i % qs.nsets
        eng.query_sig_dev(qs.nq, qs.d_q[s].data_ptr(), qs.d_qn[s].data_ptr() if qs.euclid else 0, k, flags,
                          d_idx.data_ptr(), d_score.data_ptr(), d_count.data_ptr(), env.stream)

    for i in range(warmup):
        step(i)
    env.barrier()
    eng.stats_reset(True)
    if sampler is not None:
        sampler.start()
        time.sleep(0.25)
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    env.barrier()
    t0 = time.time()
    ev0.record()
    for i in range(steps):
        step(warmup + i)
    ev1.record()
    env.barrier()
    t1 = time.time()
    ms = env.max_over_ranks(ev0.elapsed_time(ev1))
    clocks = sampler.stop(t0, t1) if sampler is not None else None
    return ms, eng.stats(), clocks


def regime(env, method, hash_num, local_rows, nq, steps, ms, st, popc_variant=3):
    """Both roofline fractions of a timed region from the library's per-launch CUDA events."""
    euclid = method == "euclid_lsh"
    w64 = (hash_num + 63) // 64
    row_bytes = w64 * 8 + (4 if euclid else 0)
    passes = max(1, st["scan_passes"])
    qp_eff = nq * steps / float(passes)
    pass_ms = st["scan_ms"] / passes
    gbps = local_rows * row_bytes / (pass_ms * 1e-3) / 1e9
    pairs = local_rows * qp_eff / (pass_ms * 1e-3)
    out = {"queries_per_pass": qp_eff, "value": steps * nq / (ms * 1e-3), "unit": UNIT, "ms_per_step": ms / steps,
           "pass_ms": pass_ms, "hbm_GBps": gbps, "hbm_frac": gbps / env.hbm_peak,
           "scan_share_of_step": st["scan_ms"] / ms,
           "int_popc32_per_s": pairs * (hash_num / 32.0)}
    peak = env.micro.get("popc_only_per_s")
    if peak:
        out["popc_issue_frac"] = pairs * issued_popc_per_pair(hash_num, popc_variant) / peak
        out["int_frac_of_measured_popc_peak"] = out["int_popc32_per_s"] / peak
        out["binding"] = "hbm" if out["hbm_frac"] >= out["popc_issue_frac"] else "popc issue (XU pipe)"
    return out


def bench_scan(env, a, name, method, hash_num, rows, nq, k, qp, flags, steps, warmup, main, second_qp=None, second_nq=None):
    """One configuration: table build, device-timed steps, (main: throughput regime, e2e), parity."""
    torch = env.torch
    from jubatus_b200 import Engine, dist as jd, workload
    G, rank = env.world, env.rank
    euclid, minhash = method == "euclid_lsh", method == "minhash"
    w64 = (hash_num + 63) // 64
    row_bytes = w64 * 8 + (4 if euclid else 0)
    opts = {}
    if qp:
        opts["queries_per_pass"] = qp
    if main and a.popc_variant >= 0:
        opts["popc_variant"] = a.popc_variant
    if main and a.scan_variant:
        opts["scan_variant"] = a.scan_variant
    if main:
        for kv in a.opt:
            opts[kv.split("=")[0]] = int(kv.split("=")[1])
    eng = Engine(method, hash_num, env.local_rank, rank, G, **opts)
    if G > 1:
        jd.comm_init_from_torch(eng)          # the library's own NCCL communicator; queries are collective from here on
    eng.set_option("reserve_rows", rows)
    seed = workload.MASTER_SEED
    t_build = build_table(env, eng, method, rows, seed)
    local_rows = eng.size()[1]
    assert local_rows == ((rows - rank + G - 1) // G if rows > rank else 0)   # (a shard knows the table up to its own last row)
    nsets = 4 if nq <= 256 else 2
    qs = QuerySets(env, eng, method, hash_num, rows, nq, nsets, seed)
    outs = (torch.zeros((nq, k), dtype=torch.int64, device=env.dev), torch.zeros((nq, k), dtype=torch.float32, device=env.dev),
            torch.zeros(nq, dtype=torch.int32, device=env.dev))

    sampler = ClockSampler(env.local_rank) if (main and rank == 0) else None
    ms, st, clocks = time_device_steps(env, eng, qs, k, flags, steps, warmup, outs, sampler)
    first = regime(env, method, hash_num, local_rows, nq, steps, ms, st)
    res = {"workload": workload_name(method, hash_num, rows, nq, k) + (", similar_row" if flags & 1 else ""),
           "value": first["value"], "unit": UNIT, "ms_per_step": ms / steps, "steps": steps,
           "table_build_s": round(t_build, 2), "rows_per_gpu": local_rows,
           "regimes": [first], "gpu_launches": st["kernel_launches"]}

    # the answers to set 0, for the parity leg
    eng.query_sig_dev(nq, qs.d_q[0].data_ptr(), qs.d_qn[0].data_ptr() if euclid else 0, k, flags, outs[0].data_ptr(),
                      outs[1].data_ptr(), outs[2].data_ptr(), env.stream)
    torch.cuda.synchronize()
    got = (outs[0].cpu().numpy().view(np.uint64), outs[1].cpu().numpy(), outs[2].cpu().numpy().view(np.uint32))
    assert int(got[2].min()) == min(k, rows)

    # ---- second regime on the same table (popcount-bound when the first one is HBM-bound) ----
    if second_qp is not None:
        eng.set_option("queries_per_pass", second_qp)
        if main and a.scan_variant:
            eng.set_option("scan_variant", 0)
        qs2, outs2 = qs, outs
        if second_nq and second_nq != nq:
            qs2 = QuerySets(env, eng, method, hash_num, rows, second_nq, 2, seed ^ 0x77)
            outs2 = (torch.zeros((second_nq, k), dtype=torch.int64, device=env.dev),
                     torch.zeros((second_nq, k), dtype=torch.float32, device=env.dev),
                     torch.zeros(second_nq, dtype=torch.int32, device=env.dev))
        t_steps = max(2, steps // 2)
        ms2, st2, _ = time_device_steps(env, eng, qs2, k, flags, t_steps, 2, outs2)
        res["regimes"].append(regime(env, method, hash_num, local_rows, qs2.nq, t_steps, ms2, st2))
        if qs2 is qs:   # same queries: the two regimes must give the same answers
            eng.query_sig_dev(nq, qs.d_q[0].data_ptr(), qs.d_qn[0].data_ptr() if euclid else 0, k, flags, outs[0].data_ptr(),
                              outs[1].data_ptr(), outs[2].data_ptr(), env.stream)
            torch.cuda.synchronize()
            assert np.array_equal(outs[0].cpu().numpy().view(np.uint64), got[0]), "regimes disagree"
        eng.set_option("queries_per_pass", qp)
        if main and a.scan_variant:
            eng.set_option("scan_variant", a.scan_variant)

    # ---- end to end: the host-buffer call (pinned CSR in, ids / scores out; hashing, copies and, on
    # N > 1, the collective inside) ----
    e2e = None
    if main or minhash:
        h_idx = torch.zeros((nq, k), dtype=torch.int64).pin_memory()
        h_score = torch.zeros((nq, k), dtype=torch.float32).pin_memory()
        h_count = torch.zeros(nq, dtype=torch.int32).pin_memory()

        def step_e2e(i):
            c = qs.h[i % qs.nsets]
            eng.query_csr_ptr(nq, c["indptr"].data_ptr(), 0 if minhash else c["seed"].data_ptr(),
                              c["key"].data_ptr() if minhash else 0, c["vals"].data_ptr(), k, flags, h_idx.data_ptr(),
                              h_score.data_ptr(), h_count.data_ptr())

        e2e_steps = max(2, steps // 2)
        for i in range(2):
            step_e2e(i)
        env.barrier()
        t0 = time.perf_counter()
        for i in range(e2e_steps):
            step_e2e(i)
        env.barrier()
        e2e_s = env.max_over_ranks(time.perf_counter() - t0)
        e2e = {"value": e2e_steps * nq / e2e_s, "unit": UNIT, "h2d_bytes_per_step": qs.h2d_bytes,
               "d2h_bytes_per_step": nq * k * 12 + nq * 4, "steps": e2e_steps,
               "api": "jubann_query_csr (host CSR of the query datums -> ids/scores)" +
                      ("; collective on the %d shard handles: scan, one packed ncclAllGather, merge inside libjubann.so" % G if G > 1 else "")}
        step_e2e(0)
        assert np.array_equal(h_idx.numpy().view(np.uint64), got[0]), "host-buffer path and device path disagree"
        assert np.array_equal(h_score.numpy().view(np.uint32), got[1].view(np.uint32))
        res["e2e"] = e2e

    # ---- parity in the same run, and the CPU baseline from the same scan ----
    parity = cpu = None
    if not a.skip_cpu:
        threads = os.cpu_count() or 1
        nparity = a.parity_queries if main else min(8, a.parity_queries)
        parity, secs, nparity = check_parity(env, eng, method, hash_num, rows, k, flags, qs, got, nparity, seed, threads)
        if rank == 0 and secs is None:
            res["parity"] = parity
        elif rank == 0:
            res["parity"] = parity
            cpu = {"value": nparity / secs, "unit": UNIT, "cores": threads, "kind": "port",
                   "sample": "%d of the batch's queries over the GPUs' own table of %d rows x %d bits (k=%d), oracle engine with "
                             "threads=%d row partitions, build -O3 -march=x86-64-v3 (hw popcnt)" % (nparity, rows, hash_num, k, threads),
                   "scan_GBps": nparity / secs * rows * row_bytes / 1e9}
            res["cpu_baseline"] = cpu
    eng.close()
    del qs, outs
    torch.cuda.empty_cache()
    return res, clocks, st


def bench_c5(env, a, rows, adds, scores, interval=512):
    """Config 5: jubaanomaly light_lof over lsh hash_num=256 through the C++ host stack.  Every rank is a
    replica holding the whole table (bulk-loaded to `rows`); then `adds` streaming add() calls per replica
    (each: set_row, two reverse-neighbour collections, a batched re-scoring, calc_score) with a MIX round
    (jubann_mix_allgather under the host's mixable) every `interval` adds when N > 1, then calc_score calls."""
    from jubatus_b200 import nn, workload
    G, rank = env.world, env.rank
    conv = {"string_filter_types": {}, "string_filter_rules": [], "num_filter_types": {}, "num_filter_rules": [],
            "string_types": {}, "string_rules": [], "num_types": {}, "num_rules": [{"key": "*", "type": "num"}]}
    cfg = json.dumps({"method": "light_lof", "converter": conv,
                      "parameter": {"nearest_neighbor_num": 10, "reverse_nearest_neighbor_num": 30, "method": "lsh",
                                    "parameter": {"hash_num": 256}}})
    an = nn.Anomaly(cfg, my_id="replica%d" % rank, device=env.local_rank)
    if G > 1:
        box = [an.comm_unique_id() if rank == 0 else None]
        env.dist.broadcast_object_list(box, src=0)
        an.comm_init(box[0], rank, G)
    t0 = time.time()
    chunk = 250_000
    for r0 in range(0, rows, chunk):
        n = min(chunk, rows - r0)
        g = env.torch.arange(r0, r0 + n, device=env.dev, dtype=env.torch.int64)
        ids, vals = workload.synth_rows_torch(workload.MASTER_SEED, g, NNZ, NFEAT_LOG2)   # the same rows on every replica
        an.add_bulk_numeric(ids.cpu().numpy().astype(np.uint32), vals.cpu().numpy(), prefix="b")
        del g, ids, vals
    t_load = time.time() - t0
    if G > 1:
        an.mix_allgather()            # the bulk load is known to every replica already: settle the clocks
    # the streamed datums: a disjoint stream, every replica its own rows
    ids, vals = workload.synth_rows(workload.QUERY_SEED, 10_000_000 + rank * (adds + scores), adds + scores, NNZ, NFEAT_LOG2)
    for i in range(3):
        an.calc_score_numeric(ids[adds + i % max(1, scores)], vals[adds + i % max(1, scores)])
    env.barrier()
    t0 = time.perf_counter()
    mix_rounds, mix_bytes, finite = 0, 0, []
    for i in range(adds):
        finite.append(an.add_numeric(ids[i], vals[i], prefix="r%d_" % rank)[1])
        if G > 1 and (i + 1) % interval == 0:
            stt = an.mix_allgather()
            mix_rounds += 1
            mix_bytes += stt["blob_bytes"] + stt["rows_sent"] * (8 + 32 + 8)
    if G > 1 and adds % interval:
        stt = an.mix_allgather(); mix_rounds += 1
        mix_bytes += stt["blob_bytes"] + stt["rows_sent"] * (8 + 32 + 8)
    env.barrier()
    t_add = env.max_over_ranks(time.perf_counter() - t0)
    t0 = time.perf_counter()
    for i in range(scores):
        finite.append(an.calc_score_numeric(ids[adds + i], vals[adds + i]))
    env.barrier()
    t_score = env.max_over_ranks(time.perf_counter() - t0)
    size = an.size()
    digest = an.table_digest()
    same = True
    if G > 1:
        all_d = [None] * G
        env.dist.all_gather_object(all_d, (size, digest))
        same = all(d == all_d[0] for d in all_d)
        if not same:
            raise RuntimeError("config 5: replicas differ after MIX: %r" % (all_d,))
    ok = [s for s in finite if s == s and abs(s) != float("inf")]
    res = {"workload": "light_lof over lsh hash_num=256 (nn=10, rnn=30), %d-row table per replica, streaming add/calc_score, "
                       "MIX every %d adds across %d replica(s)" % (rows, interval, G),
           "value": G * adds / t_add, "unit": "add/s (all replicas, MIX included)", "adds_per_replica": adds,
           "ms_per_add": t_add / max(1, adds) * 1e3, "calc_score_per_s": G * scores / t_score if scores else None,
           "bulk_load_s": round(t_load, 2), "rows_after": size, "mix_rounds": mix_rounds,
           "mix_bytes_sent_per_round": mix_bytes / mix_rounds if mix_rounds else 0,
           "replicas_identical_after_mix": bool(same), "mean_score": float(np.mean(ok)) if ok else None,
           "api": "jubatus_b200 host stack (C++ light_lof / mixable_versioned_table) over the C-ABI; MIX = jubann_mix_allgather + jubann_mix_apply"}
    expect = rows + G * adds
    if size != expect:
        raise RuntimeError("config 5: %d rows after the run, expected %d" % (size, expect))
    an.close()
    return res


def run_b200(a):
    env = Env(a)
    torch = env.torch
    G, rank = env.world, env.rank
    hres, clocks, st = bench_scan(env, a, "headline", a.method, a.hash_num, a.rows, a.nq, a.k, a.qp, 0, a.steps, a.warmup,
                                  main=True, second_qp=0 if a.qp else None)
    head = hres["regimes"][0]
    w64 = (a.hash_num + 63) // 64
    euclid = a.method == "euclid_lsh"
    row_bytes = w64 * 8 + (4 if euclid else 0)
    local_rows = hres["rows_per_gpu"]
    passes_per_launch = st["scan_passes"] / float(max(1, st["scan_launches"]))
    bytes_per_launch = local_rows * row_bytes * passes_per_launch
    avg_ms = st["scan_ms"] / max(1, st["scan_launches"])
    achieved = bytes_per_launch / (avg_ms * 1e-3) / 1e9
    traffic, traffic_source = None, None
    try:
        prof = json.load(open(os.path.join(ROOT, "profiles", "scan_traffic.json")))
        if prof.get("algorithmic_bytes_per_pass") == local_rows * row_bytes:   # captured on this very workload
            traffic = prof["dram_bytes_per_pass"] * passes_per_launch
            traffic_source = "profiles/scan_traffic.json: dram__bytes_read+write of one ncu --set full capture of this kernel on this " \
                             "workload (per pass, times the passes per launch); not measured in this run"
    except Exception:
        pass
    roofline = {"bound": "hbm", "achieved": achieved, "peak": env.hbm_peak, "unit": "GB/s", "frac": achieved / env.hbm_peak,
                "traffic": traffic, "traffic_source": traffic_source,
                "peak_source": "MEASURED_PEAKS.json hbm_gbs (of measured)" if env.peaks else "fallback 6650 (of fallback)",
                "kernel": scan_kernel_name(a.hash_num, head["queries_per_pass"], local_rows),
                "bytes_per_launch": bytes_per_launch, "avg_launch_ms": avg_ms, "launches": st["scan_launches"],
                "passes_per_launch": passes_per_launch, "queries_per_pass": head["queries_per_pass"],
                "scan_share_of_step": head["scan_share_of_step"], "int_popc32_per_s": head["int_popc32_per_s"],
                "popc_issue_frac": head.get("popc_issue_frac"),
                "int_frac_of_measured_popc_peak": head.get("int_frac_of_measured_popc_peak")}
    line = {
        "metric": METRIC, "value": hres["value"], "unit": UNIT, "n_gpus": G, "steps": a.steps, "warmup": a.warmup,
        "ms_per_step": hres["ms_per_step"], "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
        "dtype": "u32",
        "data": "synthetic sparse datums of SURVEY 8(d) (splitmix64 seed 0x4A554241, 100 distinct ascending ids of 2^20, values 1+U[0,1)) "
                "hashed by the signature kernel; queries from the disjoint stream with 1% exact + 1% near duplicates",
        "config": {"workload": hres["workload"], "parallelism": "rows sharded round-robin over %d GPU(s)" % G,
                   "cache": "table %.1f GB per GPU >> 126 MB L2, streamed once per pass (inputs larger than L2)" % (local_rows * row_bytes / 1e9),
                   "queries_per_pass": head["queries_per_pass"], "table_build_s": hres["table_build_s"]},
        "roofline": roofline, "e2e": hres.get("e2e"), "gpu_launches": hres["gpu_launches"], "clocks": clocks,
        "throughput_mode": hres["regimes"][1] if len(hres["regimes"]) > 1 else None,
        "parity": hres.get("parity"),
    }
    if G == 1 and "cpu_baseline" in hres:
        line["cpu_baseline"] = hres["cpu_baseline"]

    # ---- BASELINE.json's other configurations, short runs ----
    want = [c for c in a.configs.split(",") if c]
    configs = {}
    sc = a.config_scale
    short = max(2, min(4, a.steps))
    if "c2" in want:   # lsh 512 bits, 10^7 rows, 4096-query batches, k=10 (popcount regime)
        r, _, _ = bench_scan(env, a, "c2", "lsh", 512, int(10_000_000 * sc), 4096, 10, 0, 0, short, 3, main=False)
        configs["c2"] = r
    if "c3" in want:   # euclid_lsh 256 bits + norm, 10^8 rows, k=100: HBM regime at 2 per pass, then 4096-query batches
        r, _, _ = bench_scan(env, a, "c3", "euclid_lsh", 256, int(100_000_000 * sc), 64, 100, 2, 0, max(short, 4), 3, main=False,
                             second_qp=0, second_nq=4096 if sc >= 0.1 else 256)
        configs["c3"] = r
    if "c4" in want:   # nearest_neighbor_recommender / minhash 128 bits, 2*10^7 rows, similar_row_from_datum k=50
        r, _, _ = bench_scan(env, a, "c4", "minhash", 128, int(20_000_000 * sc), 4096, 50, 0, 1, short, 3, main=False)
        configs["c4"] = r
    if "c5" in want:
        configs["c5"] = bench_c5(env, a, int(1_000_000 * sc), 1024, 512)
    for name, r in configs.items():
        if "regimes" in r:
            best = max(r["regimes"], key=lambda g: max(g.get("hbm_frac", 0), g.get("popc_issue_frac", 0) or 0))
            r["roofline"] = {"kernel": scan_kernel_name(int(r["workload"].split("hash_num=")[1].split(",")[0]), r["regimes"][0]["queries_per_pass"], r["rows_per_gpu"]),
                             "hbm_frac": r["regimes"][0]["hbm_frac"], "popc_issue_frac": r["regimes"][0].get("popc_issue_frac"),
                             "binding": r["regimes"][0].get("binding"), "best_fraction_any_regime": max(best.get("hbm_frac", 0), best.get("popc_issue_frac", 0) or 0)}
    if configs:
        line["configs"] = configs
    if rank == 0:
        emit(line)
    if G > 1:
        env.dist.barrier()
        env.dist.destroy_process_group()


_REAL_STDOUT = None


def emit(line):
    """The one JSON line goes to the real stdout; everything else printed meanwhile (NCCL's version
    banner, library chatter) has been diverted to stderr."""
    data = (json.dumps(line) + "\n").encode()
    if _REAL_STDOUT is None:
        sys.stdout.write(data.decode()); sys.stdout.flush()
    else:
        os.write(_REAL_STDOUT, data)


def main():
    global _REAL_STDOUT
    a = parse_args()
    sys.stdout.flush()
    _REAL_STDOUT = os.dup(1)
    os.dup2(2, 1)
    try:
        if a.impl == "reference":
            run_reference(a)
        else:
            run_b200(a)
    except BaseException:
        # a rank that fails must not linger in a destructor that waits for its peers (they would wait
        # for it in the next collective): report and leave at once, so that the launcher ends the job
        import traceback
        traceback.print_exc()
        sys.stderr.flush()
        os._exit(1)


if __name__ == "__main__":
    main()
