"""jubatus_b200 -- B200-native backend for Jubatus' nearest_neighbor engines.

The product is the C-ABI library ``libjubann.so`` (CUDA, sm_100a; declared in
``include/jubann.h``).  This package is the thin Python face of it used by the tests, the
bench and the multi-GPU plumbing (``torch.distributed`` over NCCL): ctypes bindings
(:mod:`jubatus_b200.capi`), the reference-shaped host layer (:mod:`jubatus_b200.nn`), and
the sharded / replicated multi-GPU drivers (:mod:`jubatus_b200.dist`).

There is no CPU fallback: loading fails loudly if the CUDA library has not been built.
"""
from .capi import (  # noqa: F401
    JubannError,
    Engine,
    lib_path,
    load_library,
    SIMILAR,
)

__all__ = ["JubannError", "Engine", "lib_path", "load_library", "SIMILAR"]
