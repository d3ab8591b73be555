"""Synthetic workload of SURVEY.md 8(d): sparse datums of `nnz` hashed features.

splitmix64-based, stateless, master seed 0x4A554241 ("JUBA").  Row r: the first `nnz`
distinct ids among draws t = 0, 1, ... (id = draw & (2^nfeat_log2 - 1)), sorted ascending;
value of feature id f = 1 + (draw(r, 2^32 + f) >> 40) / 2^24.  Feature f is the datum key
"<f>" under the `num` rule, i.e. the sfv key "<f>@num"; seed32 / keyhash64 are the
FNV-1 hashes of that key (what the engines derive from sfv keys).

This numpy generator is the product-side twin of oracle::synth_row (tests check they
agree); it is input generation, not part of the measured path.
"""
import numpy as np

MASTER_SEED = 0x4A554241
QUERY_SEED = MASTER_SEED ^ 0x51
_M64 = np.uint64(0xFFFFFFFFFFFFFFFF)


def splitmix64(x):
    x = np.asarray(x, dtype=np.uint64)
    with np.errstate(over="ignore"):
        x = x + np.uint64(0x9E3779B97F4A7C15)
        x = (x ^ (x >> np.uint64(30))) * np.uint64(0xBF58476D1CE4E5B9)
        x = (x ^ (x >> np.uint64(27))) * np.uint64(0x94D049BB133111EB)
        return x ^ (x >> np.uint64(31))


def synth_draw(seed, row, t):
    with np.errstate(over="ignore"):
        return splitmix64(splitmix64(np.uint64(seed) ^ splitmix64(row)) + np.asarray(t, dtype=np.uint64))


def fnv1_64(data):
    h = 14695981039346656037
    for c in data:
        h = (h * 1099511628211) & 0xFFFFFFFFFFFFFFFF
        if c >= 0x80:
            c |= 0xFFFFFFFFFFFFFF00  # `char` sign-extends in the reference
        h ^= c
    return h


def feature_tables(nfeat_log2=20):
    """seed32[f], keyhash64[f] for key '<f>@num' (vectorised FNV-1 over the digit strings)."""
    n = 1 << nfeat_log2
    f = np.arange(n, dtype=np.uint64)
    ndig = np.ones(n, dtype=np.int64)
    p = 10
    while p <= n:
        ndig += (f >= p).astype(np.int64)
        p *= 10
    maxd = int(ndig.max())
    h = np.full(n, 14695981039346656037, dtype=np.uint64)
    prime = np.uint64(1099511628211)
    with np.errstate(over="ignore"):
        for pos in range(maxd):  # most significant digit first
            active = ndig > pos
            div = np.power(np.uint64(10), np.maximum(ndig - 1 - pos, 0).astype(np.uint64))
            digit = (f // div) % np.uint64(10) + np.uint64(ord("0"))
            hn = (h * prime) ^ digit
            h = np.where(active, hn, h)
        for c in b"@num":
            h = (h * prime) ^ np.uint64(c)
    return h.astype(np.uint32), h


def synth_rows(seed, row0, n, nnz=100, nfeat_log2=20):
    """ids[n][nnz] (uint32, ascending), vals[n][nnz] (float32)."""
    mask = np.uint64((1 << nfeat_log2) - 1)
    rows = np.arange(row0, row0 + n, dtype=np.uint64)
    ids = np.zeros((n, nnz), dtype=np.uint32)
    draws = nnz + 32
    pending = np.arange(n)
    while len(pending):
        t = np.arange(draws, dtype=np.uint64)
        d = synth_draw(seed, rows[pending, None], t[None, :]) & mask
        d = d.astype(np.uint32)
        # first occurrence of every id, in draw order
        order = np.argsort(d, axis=1, kind="stable")
        sd = np.take_along_axis(d, order, axis=1)
        first = np.ones_like(sd, dtype=bool)
        first[:, 1:] = sd[:, 1:] != sd[:, :-1]
        is_first = np.zeros_like(first)
        np.put_along_axis(is_first, order, first, axis=1)
        rank = np.cumsum(is_first, axis=1)
        ok = rank[:, -1] >= nnz
        sel = is_first & (rank <= nnz)
        good = pending[ok]
        if len(good):
            picked = d[ok][sel[ok]].reshape(len(good), nnz)
            ids[good] = np.sort(picked, axis=1)
        pending = pending[~ok]
        draws *= 2
    with np.errstate(over="ignore"):
        dv = synth_draw(seed, rows[:, None], (np.uint64(1) << np.uint64(32)) + ids.astype(np.uint64))
    vals = (np.float32(1.0) + (dv >> np.uint64(40)).astype(np.float32) * np.float32(1.0 / 16777216.0)).astype(np.float32)
    return ids, vals


def to_csr(ids, vals, seed_table, keyhash_table):
    n, nnz = ids.shape
    indptr = np.arange(n + 1, dtype=np.uint64) * np.uint64(nnz)
    flat = ids.reshape(-1)
    return indptr, seed_table[flat], keyhash_table[flat], np.ascontiguousarray(vals.reshape(-1), dtype=np.float32)


def make_queries(seed, nq, stored_ids, stored_vals, nnz=100, nfeat_log2=20):
    """Queries from a disjoint stream, plus 1% exact and 1% near duplicates of stored rows."""
    ids, vals = synth_rows(seed ^ 0x51, 0, nq, nnz, nfeat_log2)
    n = stored_ids.shape[0]
    if n:
        ndup = max(1, nq // 100)
        for j in range(min(ndup, nq)):
            src = int(splitmix64(np.uint64(seed + 7 * j)) % np.uint64(n))
            ids[j], vals[j] = stored_ids[src], stored_vals[src]
        for j in range(ndup, min(2 * ndup, nq)):
            src = int(splitmix64(np.uint64(seed + 7 * j + 3)) % np.uint64(n))
            mixed = np.unique(np.concatenate([stored_ids[src][: nnz - nnz // 10], ids[j]]))[:nnz]
            if len(mixed) == nnz:
                ids[j] = mixed
    return ids, vals


def random_signatures(seed, n, w64):
    """Uniform random signature words (scan-only benchmarks; labelled as such)."""
    i = np.arange(n * w64, dtype=np.uint64)
    with np.errstate(over="ignore"):
        return splitmix64(np.uint64(seed) + i).reshape(n, w64)


# --------------------------------------------------------------------------------------
# The same generator on a torch device (bench.py builds 10^8-row tables with it; the numpy
# form above would take minutes).  int64 tensors stand for the uint64 values: add / multiply /
# xor wrap the same way, right shifts are made logical by masking.  Bit-identical to
# synth_rows (tests/test_workload.py).
# --------------------------------------------------------------------------------------
def _s64(c):
    c &= 0xFFFFFFFFFFFFFFFF
    return c - (1 << 64) if c >= (1 << 63) else c


def _lsr(x, s):
    return (x >> s) & ((1 << (64 - s)) - 1)


def splitmix64_torch(x):
    x = x + _s64(0x9E3779B97F4A7C15)
    x = (x ^ _lsr(x, 30)) * _s64(0xBF58476D1CE4E5B9)
    x = (x ^ _lsr(x, 27)) * _s64(0x94D049BB133111EB)
    return x ^ _lsr(x, 31)


def synth_draw_torch(seed, row, t):
    return splitmix64_torch(splitmix64_torch(splitmix64_torch(row) ^ _s64(seed)) + t)


def synth_rows_torch(seed, rows, nnz=100, nfeat_log2=20):
    """rows: int64 tensor of global row numbers (any order, any device).
    Returns ids int64 [n][nnz] ascending and vals float32 [n][nnz], equal to synth_rows row by row."""
    import torch
    dev = rows.device
    n = rows.shape[0]
    mask = (1 << nfeat_log2) - 1
    draws = nnz + 32
    t = torch.arange(draws, dtype=torch.int64, device=dev)
    d = synth_draw_torch(seed, rows[:, None], t[None, :]) & mask
    sd, order = torch.sort(d, dim=1, stable=True)
    first = torch.ones_like(sd, dtype=torch.bool)
    first[:, 1:] = sd[:, 1:] != sd[:, :-1]
    is_first = torch.zeros_like(first)
    is_first.scatter_(1, order, first)
    del sd, order, first
    rank = torch.cumsum(is_first, dim=1)
    ok = rank[:, -1] >= nnz
    sel = is_first & (rank <= nnz)
    del rank, is_first
    if bool(ok.all()):
        ids = d[sel].reshape(n, nnz)
    else:   # a row with more than 32 repeated draws (never seen; kept exact all the same)
        ids = torch.zeros((n, nnz), dtype=torch.int64, device=dev)
        ids[ok] = d[ok][sel[ok]].reshape(-1, nnz)
        bad = (~ok).nonzero().reshape(-1)
        for b in bad.tolist():
            one, _ = synth_rows(seed, int(rows[b].item()), 1, nnz, nfeat_log2)
            ids[b] = torch.from_numpy(one[0].astype(np.int64)).to(dev)
    del d, sel
    ids, _ = torch.sort(ids, dim=1)
    dv = synth_draw_torch(seed, rows[:, None], ids + (1 << 32))
    vals = _lsr(dv, 40).to(torch.float32) * (1.0 / 16777216.0) + 1.0
    return ids, vals


def make_queries_lazy(seed, nq, nrows, nnz=100, nfeat_log2=20, query_row0=0):
    """make_queries without the stored rows in hand: the duplicated rows are regenerated on demand
    (row r of the table is a function of (seed, r) alone).  Returns ids, vals, and for each query
    the stored row it was copied from (-1: none; the near duplicates share 90 of 100 features)."""
    ids, vals = synth_rows(seed ^ 0x51, query_row0, nq, nnz, nfeat_log2)
    exact_of = np.full(nq, -1, dtype=np.int64)
    near_of = np.full(nq, -1, dtype=np.int64)
    if nrows:
        ndup = max(1, nq // 100)
        for j in range(min(ndup, nq)):
            src = int(splitmix64(np.uint64(seed + 7 * (query_row0 + j))) % np.uint64(nrows))
            sid, sval = synth_rows(seed, src, 1, nnz, nfeat_log2)
            ids[j], vals[j] = sid[0], sval[0]
            exact_of[j] = src
        for j in range(ndup, min(2 * ndup, nq)):
            src = int(splitmix64(np.uint64(seed + 7 * (query_row0 + j) + 3)) % np.uint64(nrows))
            sid, _ = synth_rows(seed, src, 1, nnz, nfeat_log2)
            mixed = np.unique(np.concatenate([sid[0][: nnz - nnz // 10], ids[j]]))[:nnz]
            if len(mixed) == nnz:
                ids[j] = mixed
                near_of[j] = src
    return ids, vals, exact_of, near_of
