"""ctypes face of the CPU oracle (oracle/liboracle.so).  TEST INFRASTRUCTURE ONLY.

Imported by tests/, __graft_entry__.smoke() and bench.py's CPU-baseline legs; never by the
product package.  PARITY UNPINNED: see oracle/jubann_oracle.hpp.
"""
import ctypes
import os
import subprocess

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ORACLE_DIR = os.path.join(ROOT, "oracle")

METHODS = {"lsh": 0, "euclid_lsh": 1, "minhash": 2}

_vp = ctypes.c_void_p
_libs = {}


def _ptr(a):
    return None if a is None else a.ctypes.data_as(_vp)


def load(flavour="O2"):
    """flavour: "O2" (reference flags, SWAR popcount) or "v3" (-O3 x86-64-v3, hw popcnt)."""
    if flavour in _libs:
        return _libs[flavour]
    name = "liboracle.so" if flavour == "O2" else "liboracle_v3.so"
    path = os.path.join(ORACLE_DIR, name)
    if not os.path.exists(path):
        subprocess.check_call(["make", "-C", ORACLE_DIR])
    lib = ctypes.CDLL(path)
    lib.oracle_last_error.restype = ctypes.c_char_p
    lib.oracle_fnv1_64.restype = ctypes.c_uint64
    lib.oracle_fnv1_64.argtypes = [ctypes.c_char_p, ctypes.c_uint64]
    lib.oracle_mt_ints.argtypes = [ctypes.c_uint32, ctypes.c_uint64, _vp]
    lib.oracle_mt_doubles.argtypes = [ctypes.c_uint32, ctypes.c_uint64, _vp]
    lib.oracle_mt_gaussians.argtypes = [ctypes.c_uint32, ctypes.c_uint64, _vp]
    lib.oracle_mix64.argtypes = [_vp]
    lib.oracle_minhash_calc_hash.restype = ctypes.c_float
    lib.oracle_minhash_calc_hash.argtypes = [ctypes.c_uint64, ctypes.c_uint64, ctypes.c_float]
    lib.oracle_projection_column.argtypes = [ctypes.c_uint32, ctypes.c_uint32, _vp]
    lib.oracle_signatures_csr.restype = ctypes.c_int
    lib.oracle_signatures_csr.argtypes = [ctypes.c_int, ctypes.c_uint32, ctypes.c_uint64, _vp, _vp, _vp, _vp, _vp,
                                          _vp, _vp, ctypes.c_uint]
    lib.oracle_ranking_hamming.restype = ctypes.c_uint64
    lib.oracle_ranking_hamming.argtypes = [_vp, _vp, ctypes.c_uint64, ctypes.c_uint32, ctypes.c_uint64,
                                           ctypes.c_uint, _vp, _vp]
    lib.oracle_ranking_euclid.restype = ctypes.c_uint64
    lib.oracle_ranking_euclid.argtypes = [_vp, ctypes.c_float, _vp, _vp, ctypes.c_uint64, ctypes.c_uint32,
                                          ctypes.c_uint64, ctypes.c_uint, _vp, _vp]
    lib.oracle_scan_batch.restype = ctypes.c_double
    lib.oracle_scan_batch.argtypes = [ctypes.c_int, ctypes.c_uint64, _vp, _vp, _vp, _vp, ctypes.c_uint64,
                                      ctypes.c_uint32, ctypes.c_uint64, ctypes.c_uint, ctypes.c_int, _vp, _vp, _vp]
    lib.oracle_synth_rows.argtypes = [ctypes.c_uint64, ctypes.c_uint64, ctypes.c_uint64, ctypes.c_uint32,
                                      ctypes.c_uint32, _vp, _vp, ctypes.c_uint]
    lib.oracle_synth_feature_tables.argtypes = [ctypes.c_uint32, _vp, _vp]
    lib.oracle_fill_random_words.argtypes = [ctypes.c_uint64, ctypes.c_uint64, _vp, ctypes.c_uint]
    lib.oracle_nn_create.restype = _vp
    lib.oracle_nn_create.argtypes = [ctypes.c_char_p, ctypes.c_int, ctypes.c_char_p, ctypes.c_uint]
    lib.oracle_lof_create.restype = _vp
    lib.oracle_lof_create.argtypes = [ctypes.c_char_p, ctypes.c_int, ctypes.c_int, ctypes.c_int, ctypes.c_char_p]
    lib.oracle_nn_destroy.argtypes = [_vp]
    lib.oracle_nn_clear.argtypes = [_vp]
    lib.oracle_nn_size.restype = ctypes.c_uint64
    lib.oracle_nn_size.argtypes = [_vp]
    lib.oracle_nn_set_row.argtypes = [_vp, ctypes.c_char_p, _vp, _vp, ctypes.c_uint64]
    lib.oracle_nn_delete_row.argtypes = [_vp, ctypes.c_char_p]
    lib.oracle_nn_query_datum.argtypes = [_vp, _vp, _vp, ctypes.c_uint64, ctypes.c_uint64, ctypes.c_int, _vp,
                                          ctypes.c_uint64, _vp, _vp]
    lib.oracle_nn_query_id.argtypes = [_vp, ctypes.c_char_p, ctypes.c_uint64, ctypes.c_int, _vp, ctypes.c_uint64,
                                       _vp, _vp]
    lib.oracle_nn_get_signature.argtypes = [_vp, ctypes.c_char_p, _vp, _vp]
    lib.oracle_nn_get_all_rows.argtypes = [_vp, _vp, ctypes.c_uint64, _vp]
    lib.oracle_nn_mix_pair.argtypes = [_vp, _vp]
    lib.oracle_lof_calc_score_datum.restype = ctypes.c_float
    lib.oracle_lof_calc_score_datum.argtypes = [_vp, _vp, _vp, ctypes.c_uint64]
    lib.oracle_lof_calc_score_id.restype = ctypes.c_float
    lib.oracle_lof_calc_score_id.argtypes = [_vp, ctypes.c_char_p]
    lib.oracle_lof_get_params.argtypes = [_vp, ctypes.c_char_p, _vp, _vp]
    _libs[flavour] = lib
    return lib


# ---- array-level helpers -------------------------------------------------------
def set_variants(gaussian_form=-1, euclid_score=-1):
    """SURVEY A.9 alternates (0 = the frozen choice); applies to both oracle builds."""
    for fl in ("O2", "v3"):
        lib = load(fl)
        lib.oracle_set_variants.argtypes = [ctypes.c_int, ctypes.c_int]
        lib.oracle_set_variants(int(gaussian_form), int(euclid_score))


def fnv1_64(s):
    if isinstance(s, str):
        s = s.encode("utf-8")
    return int(load().oracle_fnv1_64(s, len(s)))


def mt_ints(seed, n):
    out = np.zeros(n, dtype=np.uint32)
    load().oracle_mt_ints(seed, n, _ptr(out))
    return out


def mt_doubles(seed, n):
    out = np.zeros(n, dtype=np.float64)
    load().oracle_mt_doubles(seed, n, _ptr(out))
    return out


def mt_gaussians(seed, n):
    out = np.zeros(n, dtype=np.float64)
    load().oracle_mt_gaussians(seed, n, _ptr(out))
    return out


def mix64(a, b, c):
    abc = np.array([a, b, c], dtype=np.uint64)
    load().oracle_mix64(_ptr(abc))
    return int(abc[0]), int(abc[1]), int(abc[2])


def projection_column(seed, hash_num):
    out = np.zeros(hash_num, dtype=np.float32)
    load().oracle_projection_column(seed, hash_num, _ptr(out))
    return out


def signatures_csr(method, hash_num, indptr, seed32, keyhash64, vals, want_aux=False, threads=8):
    indptr = np.ascontiguousarray(indptr, dtype=np.uint64)
    seed32 = None if seed32 is None else np.ascontiguousarray(seed32, dtype=np.uint32)
    keyhash64 = None if keyhash64 is None else np.ascontiguousarray(keyhash64, dtype=np.uint64)
    vals = np.ascontiguousarray(vals, dtype=np.float32)
    n = len(indptr) - 1
    w = (hash_num + 63) // 64
    sigs = np.zeros((n, w), dtype=np.uint64)
    norms = np.zeros(n, dtype=np.float32)
    aux = np.zeros((n, hash_num), dtype=np.float32) if want_aux else None
    rc = load().oracle_signatures_csr(METHODS[method], hash_num, n, _ptr(indptr), _ptr(seed32), _ptr(keyhash64),
                                      _ptr(vals), _ptr(sigs), _ptr(norms), _ptr(aux), threads)
    assert rc == 0, load().oracle_last_error()
    return (sigs, norms, aux) if want_aux else (sigs, norms)


def scan_batch(method, qsigs, qnorms, rows, norms, hash_num, k, threads=1, similar=False, flavour="O2"):
    """Returns (idx[nq][k], score[nq][k], count[nq], seconds)."""
    lib = load(flavour)
    w = (hash_num + 63) // 64
    qsigs = np.ascontiguousarray(qsigs, dtype=np.uint64).reshape(-1, w)
    rows = np.ascontiguousarray(rows, dtype=np.uint64).reshape(-1, w)
    nq, n = qsigs.shape[0], rows.shape[0]
    qnorms = np.zeros(nq, dtype=np.float32) if qnorms is None else np.ascontiguousarray(qnorms, dtype=np.float32)
    norms = np.zeros(max(n, 1), dtype=np.float32) if norms is None else np.ascontiguousarray(norms, dtype=np.float32)
    kk = max(k, 1)
    idx = np.zeros((nq, kk), dtype=np.uint64)
    score = np.zeros((nq, kk), dtype=np.float32)
    count = np.zeros(nq, dtype=np.uint64)
    secs = lib.oracle_scan_batch(METHODS[method], nq, _ptr(qsigs), _ptr(qnorms), _ptr(rows), _ptr(norms), n,
                                 hash_num, k, threads, 1 if similar else 0, _ptr(idx), _ptr(score), _ptr(count))
    return idx[:, :k], score[:, :k], count, secs


def synth_rows(seed, row0, n, nnz=100, nfeat_log2=20, threads=8):
    ids = np.zeros((n, nnz), dtype=np.uint32)
    vals = np.zeros((n, nnz), dtype=np.float32)
    load().oracle_synth_rows(seed, row0, n, nnz, nfeat_log2, _ptr(ids), _ptr(vals), threads)
    return ids, vals


def synth_feature_tables(nfeat):
    seed32 = np.zeros(nfeat, dtype=np.uint32)
    keyhash = np.zeros(nfeat, dtype=np.uint64)
    load().oracle_synth_feature_tables(nfeat, _ptr(seed32), _ptr(keyhash))
    return seed32, keyhash


def fill_random_words(seed, nwords, threads=8, out=None):
    if out is None:
        out = np.zeros(nwords, dtype=np.uint64)
    load().oracle_fill_random_words(seed, nwords, _ptr(out), threads)
    return out


# ---- engine-level (string ids) ----------------------------------------------------
def _keys_array(keys):
    arr = (ctypes.c_char_p * len(keys))()
    for i, k in enumerate(keys):
        arr[i] = k.encode("utf-8") if isinstance(k, str) else k
    return arr


class OracleNN(object):
    """Oracle engine with the driver/RPC surface names (set_row, neighbor_row_from_*, ...)."""

    def __init__(self, method, hash_num, my_id="", threads=1, lof=None):
        self._lib = load()
        if lof is None:
            self._h = self._lib.oracle_nn_create(method.encode(), hash_num, my_id.encode(), threads)
        else:
            self._h = self._lib.oracle_lof_create(method.encode(), hash_num, lof[0], lof[1], my_id.encode())
        if not self._h:
            raise ValueError(self._lib.oracle_last_error().decode())
        self.hash_num = hash_num

    def __del__(self):
        if getattr(self, "_h", None):
            self._lib.oracle_nn_destroy(self._h)
            self._h = None

    def clear(self):
        self._lib.oracle_nn_clear(self._h)

    def size(self):
        return int(self._lib.oracle_nn_size(self._h))

    def set_row(self, row_id, sfv):
        keys = _keys_array([k for k, _ in sfv])
        vals = np.array([v for _, v in sfv], dtype=np.float32)
        rc = self._lib.oracle_nn_set_row(self._h, row_id.encode(), keys, _ptr(vals), len(sfv))
        assert rc == 0, self._lib.oracle_last_error()
        return True

    def delete_row(self, row_id):
        return bool(self._lib.oracle_nn_delete_row(self._h, row_id.encode()))

    def _collect(self, buf, scores, count):
        n = int(count.value)
        ids = buf.value.decode().split("\n")[:n]
        return [(ids[i], float(scores[i])) for i in range(n)]

    def _query_datum(self, sfv, k, similar):
        keys = _keys_array([kk for kk, _ in sfv])
        vals = np.array([v for _, v in sfv], dtype=np.float32)
        cap = 64 * (k + 1) + 64
        buf = ctypes.create_string_buffer(cap)
        scores = np.zeros(max(k, 1), dtype=np.float32)
        count = ctypes.c_uint64()
        rc = self._lib.oracle_nn_query_datum(self._h, keys, _ptr(vals), len(sfv), k, similar, buf, cap, _ptr(scores),
                                             ctypes.byref(count))
        assert rc == 0, self._lib.oracle_last_error()
        return self._collect(buf, scores, count)

    def _query_id(self, row_id, k, similar):
        cap = 64 * (k + 1) + 64
        buf = ctypes.create_string_buffer(cap)
        scores = np.zeros(max(k, 1), dtype=np.float32)
        count = ctypes.c_uint64()
        rc = self._lib.oracle_nn_query_id(self._h, row_id.encode(), k, similar, buf, cap, _ptr(scores),
                                          ctypes.byref(count))
        assert rc == 0, self._lib.oracle_last_error()
        return self._collect(buf, scores, count)

    def neighbor_row_from_datum(self, sfv, k):
        return self._query_datum(sfv, k, 0)

    def similar_row_from_datum(self, sfv, k):
        return self._query_datum(sfv, k, 1)

    def neighbor_row_from_id(self, row_id, k):
        return self._query_id(row_id, k, 0)

    def similar_row_from_id(self, row_id, k):
        return self._query_id(row_id, k, 1)

    def get_signature(self, row_id):
        sig = np.zeros((self.hash_num + 63) // 64, dtype=np.uint64)
        norm = ctypes.c_float()
        ok = self._lib.oracle_nn_get_signature(self._h, row_id.encode(), _ptr(sig), ctypes.byref(norm))
        return (sig, norm.value) if ok else None

    def get_all_rows(self):
        cap = 64 * (self.size() + 1) + 64
        buf = ctypes.create_string_buffer(cap)
        count = ctypes.c_uint64()
        self._lib.oracle_nn_get_all_rows(self._h, buf, cap, ctypes.byref(count))
        return buf.value.decode().split("\n")[:int(count.value)]

    def mix_with(self, other):
        self._lib.oracle_nn_mix_pair(self._h, other._h)

    # light_lof
    def calc_score(self, sfv):
        keys = _keys_array([kk for kk, _ in sfv])
        vals = np.array([v for _, v in sfv], dtype=np.float32)
        return float(self._lib.oracle_lof_calc_score_datum(self._h, keys, _ptr(vals), len(sfv)))

    def calc_score_id(self, row_id):
        return float(self._lib.oracle_lof_calc_score_id(self._h, row_id.encode()))

    def lof_params(self, row_id):
        kd, lrd = ctypes.c_float(), ctypes.c_float()
        ok = self._lib.oracle_lof_get_params(self._h, row_id.encode(), ctypes.byref(kd), ctypes.byref(lrd))
        return (kd.value, lrd.value) if ok else None
