"""ctypes bindings of libjubann.so (include/jubann.h).

Host-buffer entry points take numpy arrays; ``*_dev`` entry points take raw device
pointers (ints, e.g. ``tensor.data_ptr()``) and a CUDA stream handle.
"""
import ctypes
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
SIMILAR = 1

EXPORTS = [
    "version", "jubann_abi_version", "jubann_last_error", "jubann_create", "jubann_destroy",
    "jubann_set_option", "jubann_clear", "jubann_size", "jubann_hash_csr", "jubann_set_rows_csr",
    "jubann_set_rows_sig", "jubann_get_rows_sig", "jubann_move_row", "jubann_truncate",
    "jubann_query_sig", "jubann_query_csr", "jubann_query_rows", "jubann_query_sig_dev",
    "jubann_set_rows_sig_dev", "jubann_scan_local_dev", "jubann_merge_dev",
    "jubann_diff_record_bytes", "jubann_diff_pack_dev", "jubann_diff_apply_dev", "jubann_stats",
    "jubann_stats_reset", "jubann_query_sig_noalloc_dev", "jubann_hash_csr_dev", "jubann_set_rows_csr_dev",
    "jubann_create_multi", "jubann_comm_unique_id", "jubann_comm_init", "jubann_comm_info",
    "jubann_mix_allgather", "jubann_mix_apply", "jubann_set_rows_csr_strided_dev", "jubann_set_row_query_csr",
]
COMM_ID_BYTES = 128


class JubannError(RuntimeError):
    def __init__(self, code, message):
        super().__init__("jubann error %d: %s" % (code, message))
        self.code = code
        self.message = message


def lib_path():
    return os.environ.get("JUBANN_LIBRARY") or os.path.join(_HERE, "libjubann.so")


_lib = None

_u64p = ctypes.POINTER(ctypes.c_uint64)
_u32p = ctypes.POINTER(ctypes.c_uint32)
_f32p = ctypes.POINTER(ctypes.c_float)
_vp = ctypes.c_void_p


def _point_at_torch_nccl():
    """libjubann.so opens NCCL by SONAME on first use.  In a Python process torch brings its own build of
    libnccl.so.2 (the nvidia-nccl wheel); whichever copy is loaded first serves both, and torch does not
    import against an older system copy.  So name torch's copy to the library up front (no torch import
    needed); a C++ host has no such second copy and uses the system one."""
    if os.environ.get("JUBANN_NCCL_LIBRARY"):
        return
    try:
        import importlib.util
        spec = importlib.util.find_spec("nvidia.nccl")
        for d in (list(spec.submodule_search_locations) if spec and spec.submodule_search_locations else []):
            cand = os.path.join(d, "lib", "libnccl.so.2")
            if os.path.exists(cand):
                os.environ["JUBANN_NCCL_LIBRARY"] = cand
                return
    except Exception:
        pass


def load_library():
    """dlopen the CUDA library.  No fallback: a missing library is an error."""
    global _lib
    if _lib is not None:
        return _lib
    path = lib_path()
    if not os.path.exists(path):
        raise ImportError(
            "%s not found: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
            "(jubatus_b200 has no CPU fallback)" % path)
    _point_at_torch_nccl()
    lib = ctypes.CDLL(path)
    lib.version.restype = ctypes.c_char_p
    lib.jubann_last_error.restype = ctypes.c_char_p
    lib.jubann_abi_version.restype = ctypes.c_int
    H = _vp
    sig = {
        "jubann_create": [ctypes.c_char_p, ctypes.c_uint32, ctypes.c_int, ctypes.c_uint32, ctypes.c_uint32,
                          ctypes.POINTER(_vp)],
        "jubann_destroy": [H],
        "jubann_set_option": [H, ctypes.c_char_p, ctypes.c_int64],
        "jubann_clear": [H],
        "jubann_size": [H, _u64p, _u64p],
        "jubann_hash_csr": [H, ctypes.c_uint64, _vp, _vp, _vp, _vp, _vp, _vp],
        "jubann_set_rows_csr": [H, ctypes.c_uint64, _vp, _vp, _vp, _vp, _vp],
        "jubann_set_rows_sig": [H, ctypes.c_uint64, _vp, _vp, _vp],
        "jubann_get_rows_sig": [H, ctypes.c_uint64, _vp, _vp, _vp],
        "jubann_move_row": [H, ctypes.c_uint64, ctypes.c_uint64],
        "jubann_truncate": [H, ctypes.c_uint64],
        "jubann_query_sig": [H, ctypes.c_uint64, _vp, _vp, ctypes.c_uint32, ctypes.c_uint32, _vp, _vp, _vp],
        "jubann_query_csr": [H, ctypes.c_uint64, _vp, _vp, _vp, _vp, ctypes.c_uint32, ctypes.c_uint32, _vp, _vp, _vp],
        "jubann_query_rows": [H, ctypes.c_uint64, _vp, ctypes.c_uint32, ctypes.c_uint32, _vp, _vp, _vp],
        "jubann_set_row_query_csr": [H, ctypes.c_uint64, ctypes.c_uint64, _vp, _vp, _vp, ctypes.c_uint32, ctypes.c_uint32, _vp, _vp,
                                     _vp],
        "jubann_query_sig_dev": [H, ctypes.c_uint64, _vp, _vp, ctypes.c_uint32, ctypes.c_uint32, _vp, _vp, _vp, _vp],
        "jubann_query_sig_noalloc_dev": [H, ctypes.c_uint64, _vp, _vp, ctypes.c_uint32, ctypes.c_uint32, _vp, _vp, _vp,
                                         _vp],
        "jubann_set_rows_sig_dev": [H, ctypes.c_uint64, ctypes.c_uint64, _vp, _vp, _vp],
        "jubann_hash_csr_dev": [H, ctypes.c_uint64, ctypes.c_uint64, _vp, _vp, _vp, _vp, _vp, _vp, _vp],
        "jubann_set_rows_csr_dev": [H, ctypes.c_uint64, ctypes.c_uint64, ctypes.c_uint64, _vp, _vp, _vp, _vp, _vp],
        "jubann_set_rows_csr_strided_dev": [H, ctypes.c_uint64, ctypes.c_uint64, ctypes.c_uint64, ctypes.c_uint64, _vp, _vp, _vp, _vp,
                                            _vp],
        "jubann_scan_local_dev": [H, ctypes.c_uint64, _vp, _vp, ctypes.c_uint32, _vp, _vp, _vp],
        "jubann_merge_dev": [H, ctypes.c_uint64, ctypes.c_uint32, ctypes.c_uint32, _vp, _vp, _vp, ctypes.c_uint32,
                             _vp, _vp, _vp, _vp],
        "jubann_diff_record_bytes": [H, _u32p],
        "jubann_diff_pack_dev": [H, ctypes.c_uint64, _vp, _vp, _vp],
        "jubann_diff_apply_dev": [H, ctypes.c_uint64, _vp, _vp],
        "jubann_stats": [H, _u64p, _u64p, _u64p, ctypes.POINTER(ctypes.c_double)],
        "jubann_stats_reset": [H, ctypes.c_int],
        "jubann_create_multi": [ctypes.c_char_p, ctypes.c_uint32, ctypes.POINTER(ctypes.c_int), ctypes.c_int,
                                ctypes.POINTER(_vp)],
        "jubann_comm_unique_id": [_vp],
        "jubann_comm_init": [H, _vp, ctypes.c_uint32, ctypes.c_uint32],
        "jubann_comm_info": [H, _u32p, _u32p],
        "jubann_mix_allgather": [H, ctypes.c_uint64, _vp, _vp, ctypes.c_uint64, _vp, _vp, _vp],
        "jubann_mix_apply": [H, ctypes.c_uint64, _vp, _vp, _vp],
    }
    for name, argtypes in sig.items():
        if os.environ.get("JUBANN_LIBRARY") and not hasattr(lib, name):
            continue   # (an older build named for an A/B run: it simply lacks the newer entry points)
        fn = getattr(lib, name)
        fn.argtypes = argtypes
        fn.restype = ctypes.c_int
    _lib = lib
    return lib


def _check(rc):
    if rc != 0:
        raise JubannError(rc, load_library().jubann_last_error().decode("utf-8", "replace"))


def _ptr(a):
    return None if a is None else a.ctypes.data_as(_vp)


def _arr(a, dtype):
    if a is None:
        return None
    return np.ascontiguousarray(a, dtype=dtype)


class Engine(object):
    """One device-resident table + engine (jubann_t).

    method: "lsh" | "euclid_lsh" | "minhash"; hash_num as in the JSON config
    (config/nearest_neighbor/*.json `parameter.hash_num`).
    """

    def __init__(self, method, hash_num, device=0, shard_rank=0, shard_count=1, devices=None, **options):
        self._lib = load_library()
        self._h = _vp()
        if devices is not None:   # one process, several GPUs: rows sharded over the device list
            devs = (ctypes.c_int * len(devices))(*[int(d) for d in devices])
            _check(self._lib.jubann_create_multi(method.encode(), int(hash_num), devs, len(devices),
                                                 ctypes.byref(self._h)))
            device = devices[0] if len(devices) else 0
        else:
            _check(self._lib.jubann_create(method.encode(), int(hash_num), int(device), int(shard_rank),
                                           int(shard_count), ctypes.byref(self._h)))
        self.devices = None if devices is None else [int(d) for d in devices]
        self.method = method
        self.hash_num = int(hash_num)
        self.w64 = (self.hash_num + 63) // 64
        self.device = int(device)
        self.shard_rank = int(shard_rank)
        self.shard_count = int(shard_count)
        for k, v in options.items():
            self.set_option(k, v)

    # -- lifetime ----------------------------------------------------------
    def close(self):
        if getattr(self, "_h", None) is not None and self._h.value:
            self._lib.jubann_destroy(self._h)
            self._h = _vp()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    # -- table -------------------------------------------------------------
    def set_option(self, name, value):
        _check(self._lib.jubann_set_option(self._h, name.encode(), int(value)))

    def clear(self):
        _check(self._lib.jubann_clear(self._h))

    def size(self):
        g, l = ctypes.c_uint64(), ctypes.c_uint64()
        _check(self._lib.jubann_size(self._h, ctypes.byref(g), ctypes.byref(l)))
        return g.value, l.value

    def _csr(self, indptr, seed32, keyhash64, vals):
        indptr = _arr(indptr, np.uint64)
        return indptr, _arr(seed32, np.uint32), _arr(keyhash64, np.uint64), _arr(vals, np.float32)

    def hash_csr(self, indptr, seed32, keyhash64, vals):
        indptr, seed32, keyhash64, vals = self._csr(indptr, seed32, keyhash64, vals)
        n = len(indptr) - 1
        sigs = np.zeros((n, self.w64), dtype=np.uint64)
        norms = np.zeros(n, dtype=np.float32)
        _check(self._lib.jubann_hash_csr(self._h, n, _ptr(indptr), _ptr(seed32), _ptr(keyhash64), _ptr(vals),
                                         _ptr(sigs), _ptr(norms)))
        return sigs, norms

    def set_rows_csr(self, row_index, indptr, seed32, keyhash64, vals):
        indptr, seed32, keyhash64, vals = self._csr(indptr, seed32, keyhash64, vals)
        row_index = _arr(row_index, np.uint64)
        _check(self._lib.jubann_set_rows_csr(self._h, len(row_index), _ptr(row_index), _ptr(indptr), _ptr(seed32),
                                             _ptr(keyhash64), _ptr(vals)))

    def set_rows_sig(self, row_index, sigs, norms=None):
        row_index = _arr(row_index, np.uint64)
        sigs = _arr(sigs, np.uint64)
        norms = _arr(norms, np.float32)
        _check(self._lib.jubann_set_rows_sig(self._h, len(row_index), _ptr(row_index), _ptr(sigs), _ptr(norms)))

    def get_rows_sig(self, row_index):
        row_index = _arr(row_index, np.uint64)
        n = len(row_index)
        sigs = np.zeros((n, self.w64), dtype=np.uint64)
        norms = np.zeros(n, dtype=np.float32)
        _check(self._lib.jubann_get_rows_sig(self._h, n, _ptr(row_index), _ptr(sigs), _ptr(norms)))
        return sigs, norms

    def move_row(self, src, dst):
        _check(self._lib.jubann_move_row(self._h, int(src), int(dst)))

    def truncate(self, new_global_rows):
        _check(self._lib.jubann_truncate(self._h, int(new_global_rows)))

    # -- queries -----------------------------------------------------------
    @staticmethod
    def _outs(nq, k):
        return (np.zeros((nq, k), dtype=np.uint64), np.zeros((nq, k), dtype=np.float32),
                np.zeros(nq, dtype=np.uint32))

    def query_sig(self, qsigs, qnorms, k, flags=0):
        qsigs = _arr(qsigs, np.uint64).reshape(-1, self.w64)
        qnorms = _arr(qnorms, np.float32)
        nq = qsigs.shape[0]
        idx, score, count = self._outs(nq, k)
        _check(self._lib.jubann_query_sig(self._h, nq, _ptr(qsigs), _ptr(qnorms), int(k), int(flags), _ptr(idx),
                                          _ptr(score), _ptr(count)))
        return idx, score, count

    def query_csr(self, indptr, seed32, keyhash64, vals, k, flags=0):
        indptr, seed32, keyhash64, vals = self._csr(indptr, seed32, keyhash64, vals)
        nq = len(indptr) - 1
        idx, score, count = self._outs(nq, k)
        _check(self._lib.jubann_query_csr(self._h, nq, _ptr(indptr), _ptr(seed32), _ptr(keyhash64), _ptr(vals),
                                          int(k), int(flags), _ptr(idx), _ptr(score), _ptr(count)))
        return idx, score, count

    def set_row_query_csr(self, row_index, seed32, keyhash64, vals, k, flags=0):
        """set_row of one sparse vector, then neighbor_row(id) of it, in one call."""
        vals = _arr(vals, np.float32)
        seed32 = None if seed32 is None else _arr(seed32, np.uint32)
        keyhash64 = None if keyhash64 is None else _arr(keyhash64, np.uint64)
        idx, score, count = self._outs(1, k)
        _check(self._lib.jubann_set_row_query_csr(self._h, int(row_index), len(vals), _ptr(seed32), _ptr(keyhash64), _ptr(vals),
                                                  int(k), int(flags), _ptr(idx), _ptr(score), _ptr(count)))
        return idx, score, count

    def query_rows(self, row_index, k, flags=0):
        row_index = _arr(row_index, np.uint64)
        nq = len(row_index)
        idx, score, count = self._outs(nq, k)
        _check(self._lib.jubann_query_rows(self._h, nq, _ptr(row_index), int(k), int(flags), _ptr(idx), _ptr(score),
                                           _ptr(count)))
        return idx, score, count

    # -- raw host-pointer variant (pinned buffers, end-to-end timing) --------
    def query_sig_ptr(self, nq, qsigs_ptr, qnorms_ptr, k, flags, out_idx_ptr, out_score_ptr, out_count_ptr):
        _check(self._lib.jubann_query_sig(self._h, int(nq), _vp(qsigs_ptr), _vp(qnorms_ptr) if qnorms_ptr else None,
                                          int(k), int(flags), _vp(out_idx_ptr), _vp(out_score_ptr),
                                          _vp(out_count_ptr)))

    def query_csr_ptr(self, nq, indptr_ptr, seed32_ptr, keyhash_ptr, vals_ptr, k, flags, out_idx_ptr, out_score_ptr,
                      out_count_ptr):
        _check(self._lib.jubann_query_csr(self._h, int(nq), _vp(indptr_ptr), _vp(seed32_ptr) if seed32_ptr else None,
                                          _vp(keyhash_ptr) if keyhash_ptr else None, _vp(vals_ptr), int(k), int(flags),
                                          _vp(out_idx_ptr), _vp(out_score_ptr), _vp(out_count_ptr)))

    # -- device-pointer entry points -----------------------------------------
    def query_sig_dev(self, nq, d_qsigs, d_qnorms, k, flags, d_out_index, d_out_score, d_out_count, stream=0):
        _check(self._lib.jubann_query_sig_dev(self._h, int(nq), _vp(d_qsigs), _vp(d_qnorms) if d_qnorms else None,
                                              int(k), int(flags), _vp(d_out_index), _vp(d_out_score),
                                              _vp(d_out_count), _vp(stream) if stream else None))

    def query_sig_noalloc_dev(self, nq, d_qsigs, d_qnorms, k, flags, d_out_index, d_out_score, d_out_count, stream=0):
        """Like query_sig_dev, for callers that may not allocate: the kernels' scratch is the
        fixed 32 MiB array inside the library, so this call never reaches cudaMalloc.  Requests
        that do not fit are refused (JubannError).  Calls share that array: keep them on one stream."""
        _check(self._lib.jubann_query_sig_noalloc_dev(self._h, int(nq), _vp(d_qsigs),
                                                      _vp(d_qnorms) if d_qnorms else None, int(k), int(flags),
                                                      _vp(d_out_index), _vp(d_out_score), _vp(d_out_count),
                                                      _vp(stream) if stream else None))

    def set_rows_sig_dev(self, n, first_row_index, d_sigs, d_norms=0, stream=0):
        _check(self._lib.jubann_set_rows_sig_dev(self._h, int(n), int(first_row_index), _vp(d_sigs),
                                                 _vp(d_norms) if d_norms else None, _vp(stream) if stream else None))

    def hash_csr_dev(self, n, nnz, d_indptr, d_seed32, d_keyhash64, d_vals, d_out_sigs, d_out_norms, stream=0):
        _check(self._lib.jubann_hash_csr_dev(self._h, int(n), int(nnz), _vp(d_indptr), _vp(d_seed32) if d_seed32 else None,
                                             _vp(d_keyhash64) if d_keyhash64 else None, _vp(d_vals), _vp(d_out_sigs),
                                             _vp(d_out_norms), _vp(stream) if stream else None))

    def set_rows_csr_dev(self, n, first_row_index, nnz, d_indptr, d_seed32, d_keyhash64, d_vals, stream=0):
        _check(self._lib.jubann_set_rows_csr_dev(self._h, int(n), int(first_row_index), int(nnz), _vp(d_indptr),
                                                 _vp(d_seed32) if d_seed32 else None,
                                                 _vp(d_keyhash64) if d_keyhash64 else None, _vp(d_vals),
                                                 _vp(stream) if stream else None))

    def set_rows_csr_strided_dev(self, n, first_row_index, row_stride, nnz, d_indptr, d_seed32, d_keyhash64, d_vals, stream=0):
        _check(self._lib.jubann_set_rows_csr_strided_dev(self._h, int(n), int(first_row_index), int(row_stride), int(nnz),
                                                         _vp(d_indptr), _vp(d_seed32) if d_seed32 else None,
                                                         _vp(d_keyhash64) if d_keyhash64 else None, _vp(d_vals),
                                                         _vp(stream) if stream else None))

    def scan_local_dev(self, nq, d_qsigs, d_qnorms, k, d_cand_key, d_cand_index, stream=0):
        _check(self._lib.jubann_scan_local_dev(self._h, int(nq), _vp(d_qsigs), _vp(d_qnorms) if d_qnorms else None,
                                               int(k), _vp(d_cand_key), _vp(d_cand_index),
                                               _vp(stream) if stream else None))

    def merge_dev(self, nq, k, parts, d_cand_key, d_cand_index, d_qnorms, flags, d_out_index, d_out_score,
                  d_out_count, stream=0):
        _check(self._lib.jubann_merge_dev(self._h, int(nq), int(k), int(parts), _vp(d_cand_key), _vp(d_cand_index),
                                          _vp(d_qnorms) if d_qnorms else None, int(flags), _vp(d_out_index),
                                          _vp(d_out_score), _vp(d_out_count), _vp(stream) if stream else None))

    def diff_record_bytes(self):
        b = ctypes.c_uint32()
        _check(self._lib.jubann_diff_record_bytes(self._h, ctypes.byref(b)))
        return b.value

    def diff_pack_dev(self, n, d_row_index, d_records, stream=0):
        _check(self._lib.jubann_diff_pack_dev(self._h, int(n), _vp(d_row_index), _vp(d_records),
                                              _vp(stream) if stream else None))

    def diff_apply_dev(self, n, d_records, stream=0):
        _check(self._lib.jubann_diff_apply_dev(self._h, int(n), _vp(d_records), _vp(stream) if stream else None))

    # -- collectives under the C-ABI -------------------------------------------
    @staticmethod
    def comm_unique_id():
        """The id rank 0 hands to the other ranks (bytes of length COMM_ID_BYTES)."""
        buf = (ctypes.c_ubyte * COMM_ID_BYTES)()
        _check(load_library().jubann_comm_unique_id(ctypes.cast(buf, _vp)))
        return bytes(buf)

    def comm_init(self, comm_id, rank, world):
        buf = (ctypes.c_ubyte * COMM_ID_BYTES).from_buffer_copy(bytes(comm_id)) if comm_id is not None else None
        _check(self._lib.jubann_comm_init(self._h, ctypes.cast(buf, _vp) if buf is not None else None, int(rank), int(world)))

    def comm_info(self):
        r, w = ctypes.c_uint32(), ctypes.c_uint32()
        _check(self._lib.jubann_comm_info(self._h, ctypes.byref(r), ctypes.byref(w)))
        return r.value, w.value

    def mix_allgather(self, row_index, blob):
        """One MIX round's exchange: returns (rows per rank, [blob of every rank])."""
        _, world = self.comm_info()
        row_index = _arr(row_index, np.uint64)
        blob = bytes(blob)
        rows = np.zeros(world, dtype=np.uint64)
        sizes = np.zeros(world, dtype=np.uint64)
        ptrs = (ctypes.c_void_p * world)()
        _check(self._lib.jubann_mix_allgather(self._h, len(row_index), _ptr(row_index), ctypes.c_char_p(blob) if blob else None,
                                              len(blob), _ptr(rows), _ptr(sizes), ctypes.cast(ptrs, _vp)))
        blobs = [ctypes.string_at(ptrs[r], int(sizes[r])) if sizes[r] else b"" for r in range(world)]
        return rows, blobs

    def mix_apply(self, src_rank, src_pos, dst_row):
        src_rank = _arr(src_rank, np.uint32)
        src_pos = _arr(src_pos, np.uint32)
        dst_row = _arr(dst_row, np.uint64)
        _check(self._lib.jubann_mix_apply(self._h, len(dst_row), _ptr(src_rank), _ptr(src_pos), _ptr(dst_row)))

    # -- stats ---------------------------------------------------------------
    def stats(self):
        a, b, p, c = ctypes.c_uint64(), ctypes.c_uint64(), ctypes.c_uint64(), ctypes.c_double()
        _check(self._lib.jubann_stats(self._h, ctypes.byref(a), ctypes.byref(b), ctypes.byref(p), ctypes.byref(c)))
        return {"kernel_launches": a.value, "scan_launches": b.value, "scan_passes": p.value, "scan_ms": c.value}

    def stats_reset(self, time_scan_kernels=False):
        _check(self._lib.jubann_stats_reset(self._h, 1 if time_scan_kernels else 0))
