#!/usr/bin/env python
"""Headline benchmark: neighbor queries/sec @ 100M x 512-bit LSH rows, k=100 (BASELINE.json).

    python bench.py [--gpus N] [--steps K] [--warmup W]            # this repo's CUDA path
    python bench.py --impl reference [--gpus N] --steps K --warmup W  # CPU engine (oracle) on host cores

A step = one batch of --nq neighbor queries (k=100) against the full table.  The table
(100M rows x 512 bits = 6.4 GB; 64x the 126 MB L2, so nothing is cache-resident between
passes) is row-sharded round-robin over the N GPUs; every rank scans its
shard, the shard-local top-k candidates are all-gathered over NCCL and merged bit-exactly.
`value` times the device-resident path (CUDA events on the launching stream, max over
ranks); `e2e` times the host-buffer C-ABI call with pinned host memory, H2D and D2H inside.
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
TABLE_SEED = 0x4A554241


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
    ap.add_argument("--table", default="datums", choices=["datums", "words"],
                    help="datums: rows are synthetic sparse datums hashed by the signature kernel (end to end); "
                         "words: uniform random signature words (scan-only fill)")
    ap.add_argument("--skip-cpu", action="store_true", help="omit the cpu_baseline leg")
    ap.add_argument("--cpu-seconds", type=float, default=20.0, help="budget of the cpu_baseline sample")
    return ap.parse_args()


def workload_name(a):
    return "%s hash_num=%d, %d rows, %d-query batches, k=%d" % (a.method, a.hash_num, a.rows, a.nq, a.k)


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
def cpu_table(rows, w64, threads):
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import oracle_lib
    table = np.empty((rows, w64), dtype=np.uint64)
    oracle_lib.fill_random_words(TABLE_SEED, rows * w64, threads=threads, out=table.reshape(-1))
    return oracle_lib, table


def make_queries(table_rows_fn, nq, w64, seed):
    """nq query signatures: random words; every 8th is an exact copy of a stored row."""
    from jubatus_b200 import workload
    q = workload.random_signatures(seed, nq, w64)
    for j in range(0, nq, 8):
        q[j] = table_rows_fn(j)
    return q


def cpu_rows_that_fit(rows, w64):
    try:
        avail = int([l for l in open("/proc/meminfo") if l.startswith("MemAvailable")][0].split()[1]) * 1024
    except Exception:
        avail = 8 << 30
    limit = int(avail * 0.5) // (w64 * 8)
    return min(rows, max(1_000_000, limit))


def run_cpu_sample(a, seconds, flavours=("v3", "O2")):
    """Times the CPU engine on a bounded sample; returns the cpu_baseline object."""
    threads = os.cpu_count() or 1
    w64 = (a.hash_num + 63) // 64
    n_cpu = cpu_rows_that_fit(a.rows, w64)
    oracle_lib, table = cpu_table(n_cpu, w64, threads)
    norms = None
    qn = None
    if a.method == "euclid_lsh":
        norms = np.ones(n_cpu, dtype=np.float32)
    out = {}
    for fl in flavours:
        q1 = make_queries(lambda j: table[(j * 7919) % n_cpu], 1, w64, 99)
        if a.method == "euclid_lsh":
            qn = np.ones(1, dtype=np.float32)
        _, _, _, t1 = oracle_lib.scan_batch(a.method, q1, qn, table, norms, a.hash_num, a.k, threads=threads, flavour=fl)
        nq = int(max(2, min(64, (seconds / len(flavours)) / max(t1, 1e-4))))
        q = make_queries(lambda j: table[(j * 7919) % n_cpu], nq, w64, 7)
        if a.method == "euclid_lsh":
            qn = np.ones(nq, dtype=np.float32)
        _, _, _, secs = oracle_lib.scan_batch(a.method, q, qn, table, norms, a.hash_num, a.k, threads=threads, flavour=fl)
        qps = nq / secs * (n_cpu / float(a.rows))   # per-row normalised when the table had to shrink
        out[fl] = (qps, nq, secs)
    best = max(out, key=lambda f: out[f][0])
    sample = "%d queries over %d rows x %d bits (k=%d), oracle engine with threads=%d row partitions" % (
        out[best][1], n_cpu, a.hash_num, a.k, threads)
    if n_cpu != a.rows:
        sample += "; table shrunk to fit host RAM, value scaled by rows ratio %.3f" % (n_cpu / float(a.rows))
    sample += "; builds: " + ", ".join("%s=%.2f q/s" % ("O3-x86-64-v3(hw popcnt)" if f == "v3" else "O2(reference flags, SWAR popcount)", out[f][0]) for f in out)
    return {"value": out[best][0], "unit": UNIT, "cores": threads, "kind": "port", "sample": sample,
            "scan_GBps": out[best][0] * a.rows * (a.hash_num / 8 + (4 if a.method == "euclid_lsh" else 0)) / 1e9}


def run_reference(a):
    """--impl reference: the CPU engine (oracle port; the reference itself cannot be built here,
    DESIGN.md) with every host thread, on the same workload; each step is a bounded sample."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    threads = os.cpu_count() or 1
    w64 = (a.hash_num + 63) // 64
    n_cpu = cpu_rows_that_fit(a.rows, w64)
    oracle_lib, table = cpu_table(n_cpu, w64, threads)
    norms = np.ones(n_cpu, dtype=np.float32) if a.method == "euclid_lsh" else None
    fl = "v3"

    def step(nq, seed):
        q = make_queries(lambda j: table[(j * 7919 + seed) % n_cpu], nq, w64, seed)
        qn = np.ones(nq, dtype=np.float32) if a.method == "euclid_lsh" else None
        return oracle_lib.scan_batch(a.method, q, qn, table, norms, a.hash_num, a.k, threads=threads, flavour=fl)[3]

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
        "vs_baseline": None, "dtype": "u64", "data": "synthetic",
        "config": {"workload": workload_name(a), "note": "CPU engine = oracle port of jubatus_core 0.3.2 (reference not buildable here)"},
        "cpu_baseline": {"value": qps, "unit": UNIT, "cores": threads, "kind": "port", "sample": sample},
        "e2e": {"value": qps, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }
    emit(line)


# --------------------------------------------------------------------------------------
# CUDA arm
# --------------------------------------------------------------------------------------
def run_b200(a):
    import torch
    import torch.distributed as dist
    from jubatus_b200 import Engine, workload

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if world != a.gpus and world > 1:
        a.gpus = world
    if not torch.cuda.is_available():
        raise RuntimeError("bench.py needs a CUDA device (no CPU fallback)")
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    euclid = a.method == "euclid_lsh"
    w64 = (a.hash_num + 63) // 64
    G = world
    opts = {}
    if a.qp:
        opts["queries_per_pass"] = a.qp
    if a.popc_variant >= 0:
        opts["popc_variant"] = a.popc_variant
    if a.scan_variant:
        opts["scan_variant"] = a.scan_variant
    eng = Engine(a.method, a.hash_num, local_rank, rank, G, **opts)
    eng.set_option("reserve_rows", a.rows)
    stream = torch.cuda.current_stream().cuda_stream

    nnz = 100
    nsets = 4
    minhash = a.method == "minhash"
    if a.table == "words":
        # ---- table: uniform random signature words, written straight into the shard's slab ----
        chunk = 4_000_000
        gen = torch.Generator(device=dev)
        for r0 in range(0, a.rows, chunk):
            n = min(chunk, a.rows - r0)
            gen.manual_seed(TABLE_SEED + r0)   # every rank generates the same global rows
            sig = torch.randint(-2 ** 63, 2 ** 63 - 1, (n, w64), dtype=torch.int64, device=dev, generator=gen)
            nrm = (torch.rand(n, device=dev, generator=gen) + 0.5) if euclid else None
            eng.set_rows_sig_dev(n, r0, sig.data_ptr(), nrm.data_ptr() if euclid else 0, stream)
            if r0 == 0:
                keep_rows = sig[:4096].clone()
            torch.cuda.synchronize()
            del sig, nrm

        def queries(seed):
            q = torch.from_numpy(workload.random_signatures(seed, a.nq, w64).view(np.int64)).clone()
            q[::8] = keep_rows[(torch.arange(0, a.nq, 8) * 13) % keep_rows.shape[0]].cpu()
            return q

        h_q = [queries(1000 + i).pin_memory() for i in range(nsets)]
        d_q = [x.to(dev) for x in h_q]
        h_qn = [torch.ones(a.nq, dtype=torch.float32).pin_memory() for _ in range(nsets)] if euclid else None
        d_qn = [x.to(dev) for x in h_qn] if euclid else None
        h_csr = None
    else:
        # ---- table: synthetic sparse datums (100 of 2^20 hashed features, values 1 + U[0,1)), hashed on
        # the device by the signature kernel and stored: set_row end to end ----
        s32, k64 = workload.feature_tables(20)
        d_s32 = torch.from_numpy(s32.view(np.int32)).to(dev)
        d_k64 = torch.from_numpy(k64.view(np.int64)).to(dev)
        chunk = 1_000_000
        gen = torch.Generator(device=dev)
        t_build = time.time()
        for r0 in range(0, a.rows, chunk):
            n = min(chunk, a.rows - r0)
            gen.manual_seed(TABLE_SEED + r0)   # every rank generates (and hashes) the same global rows
            ids = torch.randint(0, 1 << 20, (n, nnz), device=dev, generator=gen)
            vals = (torch.rand((n, nnz), device=dev, generator=gen) + 1.0)
            flat = ids.reshape(-1)
            seed = None if minhash else d_s32[flat].contiguous()
            keyh = d_k64[flat].contiguous() if minhash else None
            indptr = torch.arange(n + 1, device=dev, dtype=torch.int64) * nnz
            eng.set_rows_csr_dev(n, r0, n * nnz, indptr.data_ptr(), seed.data_ptr() if seed is not None else 0,
                                 keyh.data_ptr() if keyh is not None else 0, vals.reshape(-1).data_ptr(), stream)
            if r0 == 0:
                keep_ids, keep_vals = ids[:4096].clone(), vals[:4096].clone()
            torch.cuda.synchronize()
            del ids, vals, flat, seed, keyh, indptr
        t_build = time.time() - t_build

        # queries: datums from another stream; every 8th is an exact copy of a stored datum
        def query_csr(seed_):
            g2 = torch.Generator(device=dev); g2.manual_seed(seed_)
            ids = torch.randint(0, 1 << 20, (a.nq, nnz), device=dev, generator=g2)
            vals = torch.rand((a.nq, nnz), device=dev, generator=g2) + 1.0
            src = (torch.arange(0, a.nq, 8, device=dev) * 13) % keep_ids.shape[0]
            ids[::8], vals[::8] = keep_ids[src], keep_vals[src]
            flat = ids.reshape(-1)
            return {"indptr": (torch.arange(a.nq + 1, dtype=torch.int64) * nnz).pin_memory(),
                    "seed": d_s32[flat].cpu().pin_memory(), "key": d_k64[flat].cpu().pin_memory(),
                    "vals": vals.reshape(-1).cpu().pin_memory()}

        h_csr = [query_csr(1000 + i) for i in range(nsets)]
        d_csr = [{k_: v.to(dev) for k_, v in c.items()} for c in h_csr]
        d_q = [torch.zeros((a.nq, w64), dtype=torch.int64, device=dev) for _ in range(nsets)]
        d_qn = [torch.zeros(a.nq, dtype=torch.float32, device=dev) for _ in range(nsets)]
        for i in range(nsets):   # query signatures resident in HBM for the device-timed region
            c = d_csr[i]
            eng.hash_csr_dev(a.nq, a.nq * nnz, c["indptr"].data_ptr(), 0 if minhash else c["seed"].data_ptr(),
                             c["key"].data_ptr() if minhash else 0, c["vals"].data_ptr(), d_q[i].data_ptr(),
                             d_qn[i].data_ptr(), stream)
        torch.cuda.synchronize()
        h_q = [x.cpu().pin_memory() for x in d_q]
        h_qn = [x.cpu().pin_memory() for x in d_qn] if euclid else None
        if not euclid:
            d_qn = None
    assert eng.size()[0] == a.rows

    d_idx = torch.zeros((a.nq, a.k), dtype=torch.int64, device=dev)
    d_score = torch.zeros((a.nq, a.k), dtype=torch.float32, device=dev)
    d_count = torch.zeros(a.nq, dtype=torch.int32, device=dev)
    if G > 1:
        ck = torch.empty((a.nq, a.k), dtype=torch.int32, device=dev)
        ci = torch.empty((a.nq, a.k), dtype=torch.int64, device=dev)
        all_k = torch.empty((G, a.nq, a.k), dtype=torch.int32, device=dev)
        all_i = torch.empty((G, a.nq, a.k), dtype=torch.int64, device=dev)

    def step_device(i):
        qn = d_qn[i % nsets].data_ptr() if euclid else 0
        if G == 1:
            eng.query_sig_dev(a.nq, d_q[i % nsets].data_ptr(), qn, a.k, 0, d_idx.data_ptr(), d_score.data_ptr(),
                              d_count.data_ptr(), stream)
        else:
            eng.scan_local_dev(a.nq, d_q[i % nsets].data_ptr(), qn, a.k, ck.data_ptr(), ci.data_ptr(), stream)
            dist.all_gather_into_tensor(all_k, ck)
            dist.all_gather_into_tensor(all_i, ci)
            eng.merge_dev(a.nq, a.k, G, all_k.data_ptr(), all_i.data_ptr(), qn, 0, d_idx.data_ptr(), d_score.data_ptr(),
                          d_count.data_ptr(), stream)

    def barrier():
        torch.cuda.synchronize()
        if G > 1:
            dist.barrier()
            torch.cuda.synchronize()

    def max_over_ranks(x):
        if G == 1:
            return x
        t = torch.tensor([x], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    # ---- device-resident timing ----
    for i in range(a.warmup):
        step_device(i)
    barrier()
    eng.stats_reset(True)
    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
        time.sleep(0.25)
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    t_wall0 = time.time()
    ev0.record()
    for i in range(a.steps):
        step_device(a.warmup + i)
    ev1.record()
    barrier()
    t_wall1 = time.time()
    ms_total = max_over_ranks(ev0.elapsed_time(ev1))
    clocks = sampler.stop(t_wall0, t_wall1) if rank == 0 else None
    st = eng.stats()
    value = a.steps * a.nq / (ms_total * 1e-3)

    # sanity: the planted exact copies come back first with distance 0
    torch.cuda.synchronize()
    last = (a.warmup + a.steps - 1) % nsets
    got_idx = d_idx.cpu().numpy()
    got_score = d_score.cpu().numpy()
    assert int(d_count.min().item()) == min(a.k, a.rows)
    if not euclid:
        planted = ((np.arange(0, a.nq, 8) * 13) % 4096)
        assert (got_score[::8, 0] == 0).all() and (got_idx[::8, 0] == planted).all(), "planted rows not found"
        assert (np.diff(got_score, axis=1) >= 0).all()

    # ---- roofline of the dominant kernel (scan), from live CUDA-event timings per launch ----
    local_rows = eng.size()[1]
    row_bytes = a.hash_num // 8 + (4 if euclid else 0) if a.hash_num % 64 == 0 else w64 * 8 + (4 if euclid else 0)
    bytes_per_launch = local_rows * row_bytes
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    peak = float(peaks.get("hbm_gbs", 6650.0))
    passes_per_launch = st["scan_passes"] / float(max(1, st["scan_launches"]))
    bytes_per_launch = bytes_per_launch * passes_per_launch   # a launch streams the shard once per pass
    avg_ms = st["scan_ms"] / max(1, st["scan_launches"])
    achieved = bytes_per_launch / (avg_ms * 1e-3) / 1e9
    traffic = None
    micro = {}
    try:
        prof = json.load(open(os.path.join(ROOT, "profiles", "scan_traffic.json")))
        if prof.get("algorithmic_bytes_per_pass") == local_rows * row_bytes:   # captured on this very workload
            traffic = prof["dram_bytes_per_pass"] * (st["scan_passes"] / float(max(1, st["scan_launches"])))
    except Exception:
        pass
    try:
        micro = json.load(open(os.path.join(ROOT, "profiles", "microbench.json")))
    except Exception:
        pass
    qp_eff = (a.nq * a.steps) / float(max(1, st["scan_passes"]))
    roofline = {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                "traffic": traffic, "peak_source": "MEASURED_PEAKS.json hbm_gbs (of measured)" if peaks else "fallback 6650 (of fallback)",
                "kernel": "scan_*_kernel", "bytes_per_launch": bytes_per_launch, "avg_launch_ms": avg_ms,
                "launches": st["scan_launches"], "passes_per_launch": passes_per_launch, "queries_per_pass": qp_eff,
                "scan_share_of_step": st["scan_ms"] / ms_total}
    popc32 = local_rows * qp_eff * passes_per_launch * (a.hash_num / 32.0) / (avg_ms * 1e-3)
    roofline["int_popc32_per_s"] = popc32
    if micro.get("popc_only_per_s"):
        roofline["int_frac_of_measured_popc_peak"] = popc32 / micro["popc_only_per_s"]

    # ---- second regime, same table: the library's default of 32 queries per pass, where the scan
    # is bound by the popcount pipes instead of HBM (SURVEY 8d asks for both fractions) ----
    throughput_mode = None
    if a.qp:
        eng.set_option("queries_per_pass", 0)
        eng.set_option("scan_variant", 0)
        for i in range(2):
            step_device(i)
        barrier()
        eng.stats_reset(True)
        t_steps = max(2, a.steps // 2)
        ev0.record()
        for i in range(t_steps):
            step_device(i)
        ev1.record()
        barrier()
        t_ms = max_over_ranks(ev0.elapsed_time(ev1))
        st2 = eng.stats()
        qp2 = (a.nq * t_steps) / float(max(1, st2["scan_passes"]))
        pass_ms = st2["scan_ms"] / max(1, st2["scan_passes"])
        popc2 = local_rows * qp2 * (a.hash_num / 32.0) / (pass_ms * 1e-3)
        throughput_mode = {"queries_per_pass": qp2, "value": t_steps * a.nq / (t_ms * 1e-3), "unit": UNIT,
                           "ms_per_step": t_ms / t_steps, "hbm_GBps": local_rows * row_bytes / (pass_ms * 1e-3) / 1e9,
                           "hbm_frac": local_rows * row_bytes / (pass_ms * 1e-3) / 1e9 / peak,
                           "int_popc32_per_s": popc2,
                           "int_frac_of_measured_popc_peak": popc2 / micro["popc_only_per_s"] if micro.get("popc_only_per_s") else None,
                           "note": "algorithmic popcounts (16 per 512-bit pair); the kernel issues 9 after carry-save adders"}
        eng.set_option("queries_per_pass", a.qp)
        if a.scan_variant:
            eng.set_option("scan_variant", a.scan_variant)

    # ---- end to end: host buffers through the C-ABI (pinned; H2D + D2H inside the call) ----
    h_idx = torch.zeros((a.nq, a.k), dtype=torch.int64).pin_memory()
    h_score = torch.zeros((a.nq, a.k), dtype=torch.float32).pin_memory()
    h_count = torch.zeros(a.nq, dtype=torch.int32).pin_memory()
    h2d = a.nq * w64 * 8 + (a.nq * 4 if euclid else 0)
    d2h = a.nq * a.k * 12 + a.nq * 4

    if h_csr is not None:
        h2d = a.nq * nnz * (4 + (8 if minhash else 4)) + (a.nq + 1) * 8

    def step_e2e(i):
        qn_h = h_qn[i % nsets].data_ptr() if euclid else 0
        if G == 1 and h_csr is not None:
            # neighbor_row_from_datum as the adapter calls it: host CSR of the query datums in, ids/scores out
            c = h_csr[i % nsets]
            eng.query_csr_ptr(a.nq, c["indptr"].data_ptr(), 0 if minhash else c["seed"].data_ptr(),
                              c["key"].data_ptr() if minhash else 0, c["vals"].data_ptr(), a.k, 0, h_idx.data_ptr(),
                              h_score.data_ptr(), h_count.data_ptr())
        elif G == 1:
            eng.query_sig_ptr(a.nq, h_q[i % nsets].data_ptr(), qn_h, a.k, 0, h_idx.data_ptr(), h_score.data_ptr(),
                              h_count.data_ptr())
        else:
            if h_csr is not None:
                c, dc = h_csr[i % nsets], d_csr[i % nsets]
                for k_ in ("indptr", "vals", "key" if minhash else "seed"):
                    dc[k_].copy_(c[k_], non_blocking=True)
                eng.hash_csr_dev(a.nq, a.nq * nnz, dc["indptr"].data_ptr(), 0 if minhash else dc["seed"].data_ptr(),
                                 dc["key"].data_ptr() if minhash else 0, dc["vals"].data_ptr(), d_q[i % nsets].data_ptr(),
                                 (d_qn[i % nsets] if euclid else scratch_norm).data_ptr(), stream)
            else:
                d_q[i % nsets].copy_(h_q[i % nsets], non_blocking=True)
                if euclid:
                    d_qn[i % nsets].copy_(h_qn[i % nsets], non_blocking=True)
            step_device(i)
            h_idx.copy_(d_idx, non_blocking=True)
            h_score.copy_(d_score, non_blocking=True)
            h_count.copy_(d_count, non_blocking=True)
            torch.cuda.synchronize()

    scratch_norm = torch.zeros(a.nq, dtype=torch.float32, device=dev)
    e2e_steps = max(2, a.steps // 2)
    for i in range(2):
        step_e2e(i)
    barrier()
    t0 = time.perf_counter()
    for i in range(e2e_steps):
        step_e2e(i)
    barrier()
    e2e_s = max_over_ranks(time.perf_counter() - t0)
    e2e = {"value": e2e_steps * a.nq / e2e_s, "unit": UNIT, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
           "steps": e2e_steps, "api": ("jubann_query_csr (host CSR of the query datums -> ids/scores)" if h_csr is not None else "jubann_query_sig (host buffers)") if G == 1 else
           "H2D + (jubann_hash_csr_dev) + jubann_scan_local_dev + NCCL all_gather + jubann_merge_dev + D2H"}
    step_device(e2e_steps - 1)          # same query set as the last end-to-end step
    torch.cuda.synchronize()
    assert np.array_equal(h_idx.numpy(), d_idx.cpu().numpy()), "host-buffer path and device path disagree"
    assert np.array_equal(h_score.numpy(), d_score.cpu().numpy())

    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": G, "steps": a.steps, "warmup": a.warmup,
        "ms_per_step": ms_total / a.steps, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
        "dtype": "u32", "data": ("synthetic sparse datums (100 of 2^20 hashed features) hashed by the signature kernel; queries are datums, "
                                 "every 8th an exact copy of a stored one" if a.table == "datums" else
                                 "synthetic (uniform random signature words, scan-only fill; queries random + planted exact copies)"),
        "config": {"workload": workload_name(a), "parallelism": "rows sharded round-robin over %d GPU(s)" % G,
                   "cache": "table %.1f GB per GPU >> 126 MB L2, streamed once per pass (inputs larger than L2)" % (local_rows * row_bytes / 1e9),
                   "queries_per_pass": qp_eff},
        "roofline": roofline, "e2e": e2e, "gpu_launches": st["kernel_launches"], "clocks": clocks,
        "throughput_mode": throughput_mode,
    }
    if G == 1 and rank == 0 and not a.skip_cpu:
        eng.close()
        del d_q
        torch.cuda.empty_cache()
        line["cpu_baseline"] = run_cpu_sample(a, a.cpu_seconds)
    if rank == 0:
        emit(line)
    if G > 1:
        dist.barrier()
        dist.destroy_process_group()


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
    if a.impl == "reference":
        run_reference(a)
    else:
        run_b200(a)


if __name__ == "__main__":
    main()
