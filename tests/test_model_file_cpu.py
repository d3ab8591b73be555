"""Model file container (jubatus/server/framework/save_load.cpp) -- no GPU needed.

Pinned by the reference: crc32 known answers of jubatus/server/common/crc32_test.cpp:26-27 and,
where oracle/_ref/libref_crc32.so exists (the reference's own crc32.cpp compiled where it lies,
oracle/Makefile target `ref`), the reference implementation itself on random buffers.  The
container layout is checked by an independent parse: struct + zlib.crc32 + the msgpack package."""
import ctypes
import json
import os
import struct
import zlib

import numpy as np
import pytest

from jubatus_b200 import nn

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CONFIG = json.dumps({"method": "lsh", "parameter": {"hash_num": 64},
                     "converter": {"num_rules": [{"key": "*", "type": "num"}]}}, indent=2)


def _ref_crc32():
    path = os.path.join(ROOT, "oracle", "_ref", "libref_crc32.so")
    if not os.path.exists(path):
        return None
    lib = ctypes.CDLL(path)
    lib.ref_calc_crc32.restype = ctypes.c_uint32
    lib.ref_calc_crc32.argtypes = [ctypes.c_char_p, ctypes.c_size_t, ctypes.c_uint32]
    return lambda data, crc=0: int(lib.ref_calc_crc32(data, len(data), crc))


def test_crc32_known_answers_of_the_reference_test():
    # crc32_test.cpp:26-27
    assert nn.crc32(b"") == 0
    assert nn.crc32(b"jubatus") == 0x41918955


def test_crc32_piecewise_like_the_reference_test():
    # crc32_test.cpp:30-52: chaining over arbitrary pieces equals one call
    rng = np.random.RandomState(7)
    data = rng.randint(0, 256, 128).astype(np.uint8).tobytes()
    want = nn.crc32(data)
    crc, i = 0, 0
    while i < len(data):
        n = min(int(rng.randint(0, 32)), len(data) - i)
        crc = nn.crc32(data[i:i + n], crc)
        i += n
    assert crc == want


def test_crc32_against_the_compiled_reference_and_zlib():
    ref = _ref_crc32()
    rng = np.random.RandomState(3)
    for n in (0, 1, 7, 28, 255, 4096, 100003):
        data = rng.randint(0, 256, n).astype(np.uint8).tobytes()
        seed = int(rng.randint(0, 2 ** 31))
        assert nn.crc32(data) == zlib.crc32(data)
        assert nn.crc32(data, seed) == zlib.crc32(data, seed)
        if ref is not None:
            assert nn.crc32(data) == ref(data) and nn.crc32(data, seed) == ref(data, seed)
    if ref is None:
        pytest.skip("oracle/_ref/libref_crc32.so not built on this machine: zlib and known answers only")


def test_container_layout_parsed_independently():
    msgpack = pytest.importorskip("msgpack")
    payload = bytes(range(256)) * 300   # 76.8 KB: raw32
    f = nn.model_write("nearest_neighbor", "model-A", CONFIG, 1, payload, timestamp=1700000123)
    assert f[:8] == b"jubatus\0"                                   # save_load.cpp:45,134
    fmt, major, minor, maint, crc, sys_n, user_n = struct.unpack(">QIIIIQQ", f[8:48])
    assert (fmt, major, minor, maint) == (1, 0, 9, 2)              # format_version; wscript:7 VERSION
    assert 48 + sys_n + user_n == len(f)
    sys_b, user_b = f[48:48 + sys_n], f[48 + sys_n:]
    # crc32 over header[0:28] + header[32:48] + system + user (save_load.cpp:84-92)
    want = zlib.crc32(user_b, zlib.crc32(sys_b, zlib.crc32(f[32:48], zlib.crc32(f[:28]))))
    assert crc == want
    assert msgpack.unpackb(sys_b, raw=False) == [1, 1700000123, "nearest_neighbor", "model-A", CONFIG]
    ver, body = msgpack.unpackb(user_b, raw=True)
    assert ver == 1 and body == payload
    assert nn.model_read(f, "nearest_neighbor", CONFIG) == payload


def test_load_checks_in_the_reference_order():
    payload = b"xyz" * 5
    f = bytearray(nn.model_write("nearest_neighbor", "m", CONFIG, 1, payload, timestamp=5))

    def reject(blob, fragment, **kw):
        with pytest.raises(nn.JubatusError) as e:
            nn.model_read(bytes(blob), kw.pop("t", "nearest_neighbor"), kw.pop("c", CONFIG), **kw)
        assert fragment in str(e.value), str(e.value)

    reject(b"jubatuX\0" + bytes(f[8:]), "invalid file format")
    reject(bytes(f[:20]), "invalid file format")
    g = bytearray(f); g[15] = 2
    reject(g, "invalid format version: 2, expected 1")
    g = bytearray(f); g[23] = 8
    reject(g, "jubatus version mismatched: 0.8.2, expected (current) version: 0.9.2")
    g = bytearray(f); g[-1] ^= 1
    reject(g, "invalid crc32 checksum")
    g = bytearray(f); g[30] ^= 0x40
    reject(g, "invalid crc32 checksum")
    reject(f, "server type mismatched: nearest_neighbor, expected recommender", t="recommender")
    reject(f, "server config mismatched", c=CONFIG.replace("64", "128"))
    reject(f, "user data version mismatched: 1, expected 2", user_data_version=2)
    # configs are compared as JSON, not as text (compare_config, save_load.cpp:102-109) ...
    reordered = json.dumps(json.loads(CONFIG), sort_keys=True, separators=(",", ":"))
    assert reordered != CONFIG
    assert nn.model_read(bytes(f), "nearest_neighbor", reordered) == payload
    # ... and not at all by load_file (overwrite_config, server_base.cpp:207-209)
    assert nn.model_read(bytes(f), "nearest_neighbor", "{}", overwrite_config=True) == payload


def test_broken_system_and_user_data_are_reported_as_such():
    msgpack = pytest.importorskip("msgpack")

    def container(sys_b, user_b):
        h = bytearray(48)
        h[:8] = b"jubatus\0"
        h[8:28] = struct.pack(">QIII", 1, 0, 9, 2)
        h[32:48] = struct.pack(">QQ", len(sys_b), len(user_b))
        crc = zlib.crc32(user_b, zlib.crc32(sys_b, zlib.crc32(bytes(h[32:48]), zlib.crc32(bytes(h[:28])))))
        h[28:32] = struct.pack(">I", crc)
        return bytes(h) + sys_b + user_b

    good_sys = msgpack.packb([1, 5, "nearest_neighbor", "m", "{}"])
    good_user = msgpack.packb([1, b"blob"])
    assert nn.model_read(container(good_sys, good_user), "nearest_neighbor", "{}") == b"blob"
    for bad_sys in (msgpack.packb([1, 5, "nearest_neighbor"]), msgpack.packb("x"), b""):
        with pytest.raises(nn.JubatusError, match="system data is broken"):
            nn.model_read(container(bad_sys, good_user), "nearest_neighbor", "{}")
    with pytest.raises(nn.JubatusError, match="invalid system data version: saved version: 2, expected 1"):
        nn.model_read(container(msgpack.packb([2, 5, "nearest_neighbor", "m", "{}"]), good_user), "nearest_neighbor", "{}")
    with pytest.raises(nn.JubatusError, match="invalid user container"):
        nn.model_read(container(good_sys, msgpack.packb([1, b"a", b"b"])), "nearest_neighbor", "{}")
    with pytest.raises(nn.JubatusError, match="user data is broken"):
        nn.model_read(container(good_sys, msgpack.packb([1, 17])), "nearest_neighbor", "{}")


def test_a_header_that_claims_huge_sections_costs_nothing():
    # sizes come from the file: a crafted header (1 TiB of user data) must fail on the short stream,
    # not by allocating what it claims
    import resource
    import struct
    header = bytearray(48)
    header[0:8] = b"jubatus\x00"
    struct.pack_into(">Q", header, 8, 1)
    struct.pack_into(">III", header, 16, 0, 9, 2)
    struct.pack_into(">Q", header, 32, 1 << 31)       # system data: 2 GiB claimed
    struct.pack_into(">Q", header, 40, 1 << 40)       # user data: 1 TiB claimed
    blob = bytes(header) + b"x" * 1000
    before = resource.getrusage(resource.RUSAGE_SELF).ru_maxrss
    with pytest.raises(nn.JubatusError) as ei:
        nn.model_read(blob, "nearest_neighbor", "{}", False, 1)
    assert "cannot read input file" in str(ei.value)
    after = resource.getrusage(resource.RUSAGE_SELF).ru_maxrss
    assert after - before < 200 * 1024                # KiB: far below the claimed sizes
