"""Reference-shaped host API over the C++ adapter (jubatus_b200/host/, libjubann_host.so).

Class and method names follow the RPC surface of the three servers that reach the
nearest-neighbor engines:

* :class:`NearestNeighbor` -- jubanearest_neighbor (jubatus/server/server/nearest_neighbor.idl:6-27)
* :class:`Recommender`     -- jubarecommender with method nearest_neighbor_recommender
  (recommender.idl:22-57; only the calls that reach the engine)
* :class:`Anomaly`         -- jubaanomaly with method light_lof (anomaly.idl:22-51)

Each takes the server's JSON config string (config/*/*.json).  All arithmetic happens in
libjubann.so on the GPU; this module only marshals strings.
"""
import ctypes
import json
import os

import numpy as np

from . import capi

_HERE = os.path.dirname(os.path.abspath(__file__))
_vp = ctypes.c_void_p
_lib = None


class JubatusError(RuntimeError):
    """A jubatus_exception from the host layer (config_exception, invalid_parameter, ...)."""


def host_lib_path():
    return os.path.join(_HERE, "libjubann_host.so")


def _load():
    global _lib
    if _lib is not None:
        return _lib
    path = host_lib_path()
    if not os.path.exists(path):
        raise ImportError("%s not found: run __graft_entry__.build()" % path)
    lib = ctypes.CDLL(path)
    lib.jubahost_last_error.restype = ctypes.c_char_p
    lib.jubahost_create.restype = _vp
    lib.jubahost_create.argtypes = [ctypes.c_char_p, ctypes.c_char_p, ctypes.c_char_p, ctypes.c_char_p, ctypes.c_int]
    lib.jubahost_create_devices.restype = _vp
    lib.jubahost_create_devices.argtypes = [ctypes.c_char_p, ctypes.c_char_p, ctypes.c_char_p, ctypes.c_char_p,
                                            ctypes.POINTER(ctypes.c_int), ctypes.c_int]
    lib.jubahost_comm_unique_id.argtypes = [_vp, _vp]
    lib.jubahost_comm_init.argtypes = [_vp, _vp, ctypes.c_uint32, ctypes.c_uint32]
    lib.jubahost_mix_allgather.argtypes = [_vp, _vp]
    lib.jubahost_destroy.argtypes = [_vp]
    lib.jubahost_convert.argtypes = [ctypes.c_char_p, _vp, _vp, ctypes.c_uint64, _vp, _vp, ctypes.c_uint64, _vp,
                                     ctypes.c_uint64, _vp, ctypes.c_uint64, _vp]
    lib.jubahost_clear.argtypes = [_vp]
    lib.jubahost_size.restype = ctypes.c_uint64
    lib.jubahost_size.argtypes = [_vp]
    lib.jubahost_set_row.argtypes = [_vp, ctypes.c_char_p, _vp, _vp, ctypes.c_uint64, _vp, _vp, ctypes.c_uint64]
    lib.jubahost_set_rows.argtypes = [_vp, ctypes.c_uint64, _vp, _vp, _vp, _vp]
    lib.jubahost_delete_row.argtypes = [_vp, ctypes.c_char_p]
    lib.jubahost_set_rows_numeric.argtypes = [_vp, ctypes.c_uint64, ctypes.c_char_p, ctypes.c_uint64, ctypes.c_uint32, _vp, _vp]
    lib.jubahost_add_numeric.argtypes = [_vp, ctypes.c_char_p, ctypes.c_uint32, _vp, _vp, _vp]
    lib.jubahost_calc_score_numeric.argtypes = [_vp, ctypes.c_uint32, _vp, _vp, _vp]
    lib.jubahost_table_digest.argtypes = [_vp, _vp]
    lib.jubahost_query_datum.argtypes = [_vp, _vp, _vp, ctypes.c_uint64, _vp, _vp, ctypes.c_uint64, ctypes.c_uint64,
                                         ctypes.c_int, _vp, ctypes.c_uint64, _vp, ctypes.c_uint64, _vp]
    lib.jubahost_query_id.argtypes = [_vp, ctypes.c_char_p, ctypes.c_uint64, ctypes.c_int, _vp, ctypes.c_uint64, _vp,
                                      ctypes.c_uint64, _vp]
    lib.jubahost_get_all_rows.argtypes = [_vp, _vp, ctypes.c_uint64, _vp]
    lib.jubahost_get_signature.argtypes = [_vp, ctypes.c_char_p, _vp, _vp]
    lib.jubahost_mix_pair.argtypes = [_vp, _vp]
    lib.jubahost_pack.restype = ctypes.c_int64
    lib.jubahost_pack.argtypes = [_vp, _vp, ctypes.c_uint64]
    lib.jubahost_unpack.argtypes = [_vp, _vp, ctypes.c_uint64]
    lib.jubahost_dump_json.argtypes = [_vp, _vp, ctypes.c_uint64]
    lib.jubahost_enable_batching.argtypes = [_vp, ctypes.c_uint64, ctypes.c_uint64]
    lib.jubahost_batch_stats.argtypes = [_vp, _vp, _vp]
    lib.jubahost_calc_score_datum.argtypes = [_vp, _vp, _vp, ctypes.c_uint64, _vp]
    lib.jubahost_calc_score_id.argtypes = [_vp, ctypes.c_char_p, _vp]
    lib.jubahost_lof_params.argtypes = [_vp, ctypes.c_char_p, _vp, _vp]
    lib.jubahost_train.argtypes = [_vp, ctypes.c_char_p, _vp, _vp, ctypes.c_uint64, _vp, _vp, ctypes.c_uint64]
    lib.jubahost_classify.argtypes = [_vp, ctypes.c_uint64, _vp, _vp, _vp, _vp, ctypes.c_uint64]
    lib.jubahost_get_labels.argtypes = [_vp, _vp, ctypes.c_uint64]
    lib.jubahost_set_label.argtypes = [_vp, ctypes.c_char_p]
    lib.jubahost_delete_label.argtypes = [_vp, ctypes.c_char_p]
    lib.jubahost_crc32.restype = ctypes.c_uint32
    lib.jubahost_crc32.argtypes = [ctypes.c_char_p, ctypes.c_uint64, ctypes.c_uint32]
    lib.jubahost_set_storage.argtypes = [_vp, ctypes.c_char_p, ctypes.c_char_p, ctypes.c_int]
    lib.jubahost_save.argtypes = [_vp, ctypes.c_char_p, _vp, ctypes.c_uint64]
    lib.jubahost_load.argtypes = [_vp, ctypes.c_char_p]
    lib.jubahost_load_file.argtypes = [_vp, ctypes.c_char_p]
    lib.jubahost_get_config.argtypes = [_vp, _vp, ctypes.c_uint64]
    lib.jubahost_model_write.restype = ctypes.c_int64
    lib.jubahost_model_write.argtypes = [ctypes.c_char_p, ctypes.c_char_p, ctypes.c_char_p, ctypes.c_uint64, ctypes.c_char_p,
                                         ctypes.c_uint64, ctypes.c_uint64, _vp, ctypes.c_uint64]
    lib.jubahost_model_read.restype = ctypes.c_int64
    lib.jubahost_model_read.argtypes = [ctypes.c_char_p, ctypes.c_uint64, ctypes.c_char_p, ctypes.c_char_p, ctypes.c_int,
                                        ctypes.c_uint64, _vp, ctypes.c_uint64]
    _lib = lib
    return lib


def _raise():
    raise JubatusError(_load().jubahost_last_error().decode("utf-8", "replace"))


def convert_weighted(converter_json, stored, query):
    """Feed `stored` [(key, string value), ...] through convert_and_update_weight (what set_row does),
    then convert `query` (key, value) the way a query does; returns its [(feature, value), ...]."""
    lib = _load()
    pairs = list(stored) + [query]
    cap = 1 << 18
    buf = ctypes.create_string_buffer(cap)
    vals = np.zeros(8192, dtype=np.float32)
    cnt = ctypes.c_uint64()
    lib.jubahost_convert_weighted.argtypes = [ctypes.c_char_p, _vp, _vp, ctypes.c_uint64, _vp, ctypes.c_uint64, _vp,
                                              ctypes.c_uint64, _vp]
    if lib.jubahost_convert_weighted(converter_json.encode(), _cstrs([k for k, _ in pairs]), _cstrs([v for _, v in pairs]),
                                     len(stored), buf, cap, vals.ctypes.data_as(_vp), len(vals), ctypes.byref(cnt)):
        _raise()
    keys = buf.value.decode("utf-8").split("\n")[:cnt.value]
    return [(k, float(vals[i])) for i, k in enumerate(keys)]


def unlearner_trace(name, param, ids):
    """Run an unlearner policy ("lru" / "random") over a sequence of touched ids ("-x" removes x);
    returns ([(position, victim), ...], final_size)."""
    lib = _load()
    buf = ctypes.create_string_buffer(1 << 20)
    size = ctypes.c_uint64()
    lib.jubahost_unlearner_trace.argtypes = [ctypes.c_char_p, ctypes.c_char_p, _vp, ctypes.c_uint64, _vp, ctypes.c_uint64, _vp]
    if lib.jubahost_unlearner_trace(name.encode(), json.dumps(param).encode(), _cstrs(ids), len(ids), buf, 1 << 20,
                                    ctypes.byref(size)):
        _raise()
    ev = [ln.split(":", 1) for ln in buf.value.decode().split("\n") if ln]
    return [(int(a), b) for a, b in ev], int(size.value)


def crc32(data, crc=0):
    """common::calc_crc32 (jubatus/server/common/crc32.cpp:25-52), chainable."""
    return int(_load().jubahost_crc32(data, len(data), crc))


def model_write(server_type, model_id, config, user_data_version, payload, timestamp=0):
    """framework::save_server's container (save_load.cpp:113-166) around `payload`; returns the file bytes."""
    lib = _load()
    args = (server_type.encode(), model_id.encode(), config.encode(), user_data_version, payload, len(payload), timestamp)
    n = lib.jubahost_model_write(*args, None, 0)
    if n < 0:
        _raise()
    buf = ctypes.create_string_buffer(int(n))
    if lib.jubahost_model_write(*args, buf, n) != n:
        _raise()
    return buf.raw


def model_read(blob, expected_type, expected_config, overwrite_config=False, user_data_version=1):
    """framework::load_server's checks (save_load.cpp:168-283); returns the driver payload."""
    lib = _load()
    args = (blob, len(blob), expected_type.encode(), expected_config.encode(), 1 if overwrite_config else 0, user_data_version)
    n = lib.jubahost_model_read(*args, None, 0)
    if n < 0:
        _raise()
    buf = ctypes.create_string_buffer(max(1, int(n)))
    if lib.jubahost_model_read(*args, buf, n) != n:
        _raise()
    return buf.raw[:n]


class Datum(object):
    """jubatus::client::common::datum (jubatus/client/common/datum.hpp:29-47)."""

    def __init__(self, values=None):
        self.string_values = []
        self.num_values = []
        if values:
            for k, v in (values.items() if isinstance(values, dict) else values):
                if isinstance(v, str):
                    self.add_string(k, v)
                else:
                    self.add_number(k, v)

    def add_string(self, key, value):
        self.string_values.append((key, value))
        return self

    def add_number(self, key, value):
        self.num_values.append((key, float(value)))
        return self


def _cstrs(items):
    arr = (ctypes.c_char_p * max(len(items), 1))()
    for i, s in enumerate(items):
        arr[i] = s.encode("utf-8") if isinstance(s, str) else s
    return arr


def _datum_args(d):
    sk = _cstrs([k for k, _ in d.string_values])
    sv = _cstrs([v for _, v in d.string_values])
    nk = _cstrs([k for k, _ in d.num_values])
    nv = np.array([v for _, v in d.num_values] or [0.0], dtype=np.float64)
    return sk, sv, len(d.string_values), nk, nv, len(d.num_values)


def convert(converter_json, d):
    """datum -> sfv with the converter section of a config (no device needed)."""
    lib = _load()
    sk, sv, ns, nk, nv, nn = _datum_args(d)
    cap = 1 << 16
    buf = ctypes.create_string_buffer(cap)
    vals = np.zeros(4096, dtype=np.float32)
    count = ctypes.c_uint64()
    rc = lib.jubahost_convert(converter_json.encode(), sk, sv, ns, nk, nv.ctypes.data_as(_vp), nn, buf, cap,
                              vals.ctypes.data_as(_vp), len(vals), ctypes.byref(count))
    if rc:
        _raise()
    keys = buf.value.decode().split("\n")[:count.value]
    return [(keys[i], float(vals[i])) for i in range(count.value)]


class _HostHandle(object):
    kind = None

    def __init__(self, config, my_id="", device=0, devices=None):
        """devices: the GPUs this server process may use; more than one shards the table over
        them inside libjubann.so (jubann_create_multi)."""
        self._lib = _load()
        capi.load_library()   # fail loudly when the CUDA library is missing
        if devices is not None:
            devs = (ctypes.c_int * len(devices))(*[int(d) for d in devices])
            self._h = self._lib.jubahost_create_devices(self.kind.encode(), config.encode(), capi.lib_path().encode(),
                                                        my_id.encode(), devs, len(devices))
        else:
            self._h = self._lib.jubahost_create(self.kind.encode(), config.encode(), capi.lib_path().encode(),
                                                my_id.encode(), int(device))
        if not self._h:
            _raise()
        self._config = config

    # MIX between server processes as collectives under the C-ABI (jubann_mix_allgather)
    def comm_unique_id(self):
        buf = (ctypes.c_ubyte * capi.COMM_ID_BYTES)()
        if self._lib.jubahost_comm_unique_id(self._h, ctypes.cast(buf, _vp)):
            _raise()
        return bytes(buf)

    def comm_init(self, comm_id, rank, world):
        buf = (ctypes.c_ubyte * capi.COMM_ID_BYTES).from_buffer_copy(bytes(comm_id)) if comm_id is not None else None
        if self._lib.jubahost_comm_init(self._h, ctypes.cast(buf, _vp) if buf is not None else None, int(rank), int(world)):
            _raise()

    def mix_allgather(self):
        """One MIX round with every replica of the communicator (all of them call it)."""
        st = (ctypes.c_uint64 * 4)()
        if self._lib.jubahost_mix_allgather(self._h, ctypes.cast(st, _vp)):
            _raise()
        return {"rows_sent": st[0], "rows_gathered": st[1], "rows_stored": st[2], "blob_bytes": st[3]}

    def table_digest(self):
        """(digest of the rows, digest of light_lof's (kdist, lrd) table): order-independent, equal on
        replicas that hold the same content."""
        out = (ctypes.c_uint64 * 2)()
        if self._lib.jubahost_table_digest(self._h, ctypes.cast(out, _vp)):
            _raise()
        return int(out[0]), int(out[1])

    def close(self):
        if getattr(self, "_h", None):
            self._lib.jubahost_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def get_config(self):
        return self._config

    def clear(self):
        if self._lib.jubahost_clear(self._h):
            _raise()
        return True

    def size(self):
        return int(self._lib.jubahost_size(self._h))

    def get_all_rows(self):
        cap = 64 * (self.size() + 1) + 64
        buf = ctypes.create_string_buffer(cap)
        count = ctypes.c_uint64()
        if self._lib.jubahost_get_all_rows(self._h, buf, cap, ctypes.byref(count)):
            _raise()
        return buf.value.decode().split("\n")[:count.value]

    def _set_row(self, row_id, d):
        sk, sv, ns, nk, nv, nn = _datum_args(d)
        if self._lib.jubahost_set_row(self._h, row_id.encode(), sk, sv, ns, nk, nv.ctypes.data_as(_vp), nn):
            _raise()
        return True

    def _query_datum(self, d, k, similar):
        sk, sv, ns, nk, nv, nn = _datum_args(d)
        cap = 64 * (k + 1) + 64
        buf = ctypes.create_string_buffer(cap)
        scores = np.zeros(max(k, 1), dtype=np.float32)
        count = ctypes.c_uint64()
        rc = self._lib.jubahost_query_datum(self._h, sk, sv, ns, nk, nv.ctypes.data_as(_vp), nn, k, similar, buf, cap,
                                            scores.ctypes.data_as(_vp), len(scores), ctypes.byref(count))
        if rc:
            _raise()
        ids = buf.value.decode().split("\n")[:count.value]
        return [(ids[i], float(scores[i])) for i in range(count.value)]

    def _query_id(self, row_id, k, similar):
        cap = 64 * (k + 1) + 64
        buf = ctypes.create_string_buffer(cap)
        scores = np.zeros(max(k, 1), dtype=np.float32)
        count = ctypes.c_uint64()
        rc = self._lib.jubahost_query_id(self._h, row_id.encode(), k, similar, buf, cap, scores.ctypes.data_as(_vp),
                                         len(scores), ctypes.byref(count))
        if rc:
            _raise()
        ids = buf.value.decode().split("\n")[:count.value]
        return [(ids[i], float(scores[i])) for i in range(count.value)]

    def get_signature(self, row_id, hash_num):
        sig = np.zeros((hash_num + 63) // 64, dtype=np.uint64)
        norm = ctypes.c_float()
        rc = self._lib.jubahost_get_signature(self._h, row_id.encode(), sig.ctypes.data_as(_vp), ctypes.byref(norm))
        if rc < 0:
            _raise()
        return None if rc else (sig, norm.value)

    def pack(self):
        """driver->pack(): the table (keys, versions, clock, columns) as one blob (save RPC payload)."""
        n = self._lib.jubahost_pack(self._h, None, 0)
        if n < 0:
            _raise()
        buf = ctypes.create_string_buffer(int(n))
        if self._lib.jubahost_pack(self._h, buf, n) != n:
            _raise()
        return buf.raw

    def unpack(self, blob):
        """driver->unpack(): replace the table by a packed one (load RPC)."""
        if self._lib.jubahost_unpack(self._h, blob, len(blob)):
            _raise()

    def get_status(self):
        """Engine part of get_status: {'data': table.dump_json()} (nearest_neighbor_serv.cpp:69-75)."""
        buf = ctypes.create_string_buffer(1024)
        if self._lib.jubahost_dump_json(self._h, buf, 1024):
            _raise()
        return {"data": buf.value.decode()}

    def mix_with(self, other):
        """One push_mixer round between two in-process servers (push_mixer.cpp:335-408)."""
        if self._lib.jubahost_mix_pair(self._h, other._h):
            _raise()


class NearestNeighbor(_HostHandle):
    kind = "nearest_neighbor"

    def set_row(self, row_id, d):
        return self._set_row(row_id, d)

    def set_rows(self, ids, datums):
        """Batched set_row (number-valued datums): one device call."""
        offs = np.zeros(len(ids) + 1, dtype=np.uint64)
        keys, vals = [], []
        for i, d in enumerate(datums):
            assert not d.string_values
            keys.extend(k for k, _ in d.num_values)
            vals.extend(v for _, v in d.num_values)
            offs[i + 1] = len(keys)
        vals = np.array(vals or [0.0], dtype=np.float64)
        rc = self._lib.jubahost_set_rows(self._h, len(ids), _cstrs(ids), offs.ctypes.data_as(_vp), _cstrs(keys),
                                         vals.ctypes.data_as(_vp))
        if rc:
            _raise()
        return True

    def neighbor_row_from_id(self, row_id, size):
        return self._query_id(row_id, size, 0)

    def neighbor_row_from_datum(self, d, size):
        return self._query_datum(d, size, 0)

    def similar_row_from_id(self, row_id, ret_num):
        return self._query_id(row_id, ret_num, 1)

    def similar_row_from_datum(self, d, ret_num):
        return self._query_datum(d, ret_num, 1)

    # ---- save / load RPCs (framework/server_base.cpp:135-212) ----
    def set_storage(self, datadir, eth="127.0.0.1", port=9199):
        if self._lib.jubahost_set_storage(self._h, datadir.encode(), eth.encode(), port):
            _raise()

    def save(self, model_id):
        """-> {"<eth>_<port>": path} like the save RPC."""
        buf = ctypes.create_string_buffer(4096)
        if self._lib.jubahost_save(self._h, model_id.encode(), buf, 4096):
            _raise()
        ident, path = buf.value.decode().split("\n", 1)
        return {ident: path}

    def load(self, model_id):
        if self._lib.jubahost_load(self._h, model_id.encode()):
            _raise()
        return True

    def load_file(self, path):
        if self._lib.jubahost_load_file(self._h, path.encode()):
            _raise()

    def get_config(self):
        buf = ctypes.create_string_buffer(1 << 16)
        if self._lib.jubahost_get_config(self._h, buf, 1 << 16):
            _raise()
        return buf.value.decode()

    def enable_batching(self, max_batch=64, max_wait_us=200):
        """Coalesce concurrent *_from_datum calls (one per RPC worker thread in the server) into
        batched device launches."""
        if self._lib.jubahost_enable_batching(self._h, int(max_batch), int(max_wait_us)):
            _raise()

    def batch_stats(self):
        b, q = ctypes.c_uint64(), ctypes.c_uint64()
        self._lib.jubahost_batch_stats(self._h, ctypes.byref(b), ctypes.byref(q))
        return {"batches": b.value, "queries": q.value}

    def delete_row(self, row_id):
        """column_table::delete_row (what an unlearner calls)."""
        rc = self._lib.jubahost_delete_row(self._h, row_id.encode())
        if rc < 0:
            _raise()
        return rc == 0


class Recommender(_HostHandle):
    kind = "recommender"

    def update_row(self, row_id, d):
        return self._set_row(row_id, d)

    def clear_row(self, row_id):
        if self._lib.jubahost_delete_row(self._h, row_id.encode()) < 0:
            _raise()
        return True

    def similar_row_from_id(self, row_id, size):
        return self._query_id(row_id, size, 1)

    def similar_row_from_datum(self, d, size):
        return self._query_datum(d, size, 1)


class Anomaly(_HostHandle):
    kind = "anomaly"

    def __init__(self, config, my_id="", device=0, devices=None):
        super().__init__(config, my_id, device, devices)
        self._next_id = 0   # global_id_generator_standalone starts at 0 (common/global_id_generator_standalone.cpp:53-57)

    def add(self, d):
        row_id = str(self._next_id)
        self._next_id += 1
        self._set_row(row_id, d)
        return row_id, self.calc_score_id(row_id)

    def update(self, row_id, d):
        self._set_row(row_id, d)
        return self.calc_score_id(row_id)

    def add_bulk(self, datums):
        """Load many number-valued datums at once (ids continue the add sequence): batched set_rows,
        then every row's kdist and lrd in batched device calls (light_lof::set_rows_bulk)."""
        ids = [str(self._next_id + i) for i in range(len(datums))]
        self._next_id += len(datums)
        offs = np.zeros(len(ids) + 1, dtype=np.uint64)
        keys, vals = [], []
        for i, d in enumerate(datums):
            keys.extend(k for k, _ in d.num_values)
            vals.extend(v for _, v in d.num_values)
            offs[i + 1] = len(keys)
        vals = np.array(vals or [0.0], dtype=np.float64)
        if self._lib.jubahost_set_rows(self._h, len(ids), _cstrs(ids), offs.ctypes.data_as(_vp), _cstrs(keys),
                                       vals.ctypes.data_as(_vp)):
            _raise()
        return ids

    # number-only datums keyed by decimal feature ids (SURVEY 8d's synthetic workload): the datum is
    # built in the C++ host layer instead of being marshalled key by key from Python
    def add_bulk_numeric(self, ids, vals, prefix=""):
        ids = np.ascontiguousarray(ids, dtype=np.uint32)
        vals = np.ascontiguousarray(vals, dtype=np.float32)
        n, nnz = ids.shape
        if self._lib.jubahost_set_rows_numeric(self._h, n, prefix.encode(), self._next_id, nnz, ids.ctypes.data_as(_vp),
                                               vals.ctypes.data_as(_vp)):
            _raise()
        first = self._next_id
        self._next_id += n
        return first

    def add_numeric(self, ids, vals, prefix=""):
        ids = np.ascontiguousarray(ids, dtype=np.uint32)
        vals = np.ascontiguousarray(vals, dtype=np.float32)
        row_id = prefix + str(self._next_id)
        self._next_id += 1
        s = ctypes.c_float()
        if self._lib.jubahost_add_numeric(self._h, row_id.encode(), len(ids), ids.ctypes.data_as(_vp), vals.ctypes.data_as(_vp),
                                          ctypes.byref(s)):
            _raise()
        return row_id, s.value

    def calc_score_numeric(self, ids, vals):
        ids = np.ascontiguousarray(ids, dtype=np.uint32)
        vals = np.ascontiguousarray(vals, dtype=np.float32)
        s = ctypes.c_float()
        if self._lib.jubahost_calc_score_numeric(self._h, len(ids), ids.ctypes.data_as(_vp), vals.ctypes.data_as(_vp),
                                                 ctypes.byref(s)):
            _raise()
        return s.value

    def calc_score(self, d):
        _, _, _, nk, nv, nn = _datum_args(d)
        s = ctypes.c_float()
        if self._lib.jubahost_calc_score_datum(self._h, nk, nv.ctypes.data_as(_vp), nn, ctypes.byref(s)):
            _raise()
        return s.value

    def calc_score_id(self, row_id):
        s = ctypes.c_float()
        if self._lib.jubahost_calc_score_id(self._h, row_id.encode(), ctypes.byref(s)):
            _raise()
        return s.value

    def lof_params(self, row_id):
        kd, lrd = ctypes.c_float(), ctypes.c_float()
        if self._lib.jubahost_lof_params(self._h, row_id.encode(), ctypes.byref(kd), ctypes.byref(lrd)):
            return None
        return kd.value, lrd.value


class Classifier(_HostHandle):
    """jubaclassifier with method "NN" (config/classifier/nn.json): train / classify / get_labels /
    set_label / delete_label / clear (classifier.idl:41-80)."""
    kind = "classifier"

    def train(self, labeled):
        """labeled: [(label, Datum), ...]; returns the number trained."""
        for label, d in labeled:
            sk, sv, ns, nk, nv, n = _datum_args(d)
            if self._lib.jubahost_train(self._h, label.encode(), sk, sv, ns, nk, nv.ctypes.data_as(_vp), n):
                _raise()
        return len(labeled)

    def classify(self, datums):
        """-> [[(label, score), ...] per datum], labels in sorted order."""
        offs = np.zeros(len(datums) + 1, dtype=np.uint64)
        keys, vals = [], []
        for i, d in enumerate(datums):
            assert not d.string_values
            keys.extend(k for k, _ in d.num_values)
            vals.extend(v for _, v in d.num_values)
            offs[i + 1] = len(keys)
        vals = np.array(vals or [0.0], dtype=np.float64)
        buf = ctypes.create_string_buffer(1 << 22)
        if self._lib.jubahost_classify(self._h, len(datums), offs.ctypes.data_as(_vp), _cstrs(keys), vals.ctypes.data_as(_vp),
                                       buf, 1 << 22):
            _raise()
        out = []
        lines = buf.value.decode().split("\n")[:len(datums)]
        for ln in lines:
            f = ln.split("\t") if ln else []
            out.append([(f[i], float(f[i + 1])) for i in range(0, len(f), 2)])
        return out

    def get_labels(self):
        buf = ctypes.create_string_buffer(1 << 20)
        if self._lib.jubahost_get_labels(self._h, buf, 1 << 20):
            _raise()
        return {a: int(b) for a, b in (ln.split("\t") for ln in buf.value.decode().split("\n") if ln)}

    def set_label(self, label):
        r = self._lib.jubahost_set_label(self._h, label.encode())
        if r < 0:
            _raise()
        return bool(r)

    def delete_label(self, label):
        r = self._lib.jubahost_delete_label(self._h, label.encode())
        if r < 0:
            _raise()
        return bool(r)
