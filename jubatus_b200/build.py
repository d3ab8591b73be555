"""Builds libjubann.so (CUDA, sm_100a) and libjubann_host.so (C++ host adapter) in-tree.

nvcc cross-compiles without a GPU.  Objects are cached under jubatus_b200/build/ and only
rebuilt when a source or header is newer.
"""
import os
import shutil
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
CSRC = os.path.join(HERE, "csrc")
HOST = os.path.join(HERE, "host")
BUILD = os.path.join(HERE, "build")
LIB = os.path.join(HERE, "libjubann.so")
HOST_LIB = os.path.join(HERE, "libjubann_host.so")

NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
    "-Xcompiler", "-fPIC", "-Xptxas", "-v",
]


# extra nvcc flags for tuning builds, e.g. JUBANN_NVCC_EXTRA=-DJBN_SCAN_TRACE (tools/scan_trace.py)
NVCC_FLAGS += [f for f in os.environ.get("JUBANN_NVCC_EXTRA", "").split() if f]


def _nvcc():
    for cand in (os.environ.get("NVCC"), shutil.which("nvcc"), "/usr/local/cuda/bin/nvcc"):
        if cand and os.path.exists(cand):
            return cand
    raise RuntimeError("nvcc not found")


def _cxx():
    """The host compiler nvcc itself uses (g++ on PATH), so that every library of the process shares
    one dynamic libstdc++.  ($CXX in this image is a wrapper that links libstdc++ statically; a
    second, older copy of the runtime inside the host library crashed in iostream code once
    libjubann.so had pulled in the system one.)"""
    return shutil.which("g++") or os.environ.get("CXX") or "g++"


def _newer(target, deps):
    if not os.path.exists(target):
        return True
    t = os.path.getmtime(target)
    return any(os.path.getmtime(d) > t for d in deps)


def _run(cmd, log=None):
    p = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
    if log:
        with open(log, "w") as f:
            f.write(" ".join(cmd) + "\n" + p.stdout)
    if p.returncode != 0:
        sys.stderr.write(p.stdout)
        raise RuntimeError("build step failed: " + " ".join(cmd))
    return p.stdout


def build_cuda(verbose=False):
    os.makedirs(BUILD, exist_ok=True)
    nvcc = _nvcc()
    headers = [os.path.join(CSRC, f) for f in os.listdir(CSRC) if f.endswith((".h", ".cuh"))]
    headers.append(os.path.join(ROOT, "include", "jubann.h"))
    sources = sorted(f for f in os.listdir(CSRC) if f.endswith(".cu"))
    objs, jobs = [], []
    for src in sources:
        s = os.path.join(CSRC, src)
        o = os.path.join(BUILD, src[:-3] + ".o")
        objs.append(o)
        if _newer(o, [s] + headers):
            jobs.append([nvcc] + NVCC_FLAGS + ["-c", s, "-o", o])
    if jobs:
        with ThreadPoolExecutor(max_workers=min(len(jobs), os.cpu_count() or 4)) as ex:
            outs = list(ex.map(lambda c: _run(c, c[-1] + ".log"), jobs))
        if verbose:
            for o in outs:
                print(o)
    if jobs or _newer(LIB, objs):
        _run([nvcc, "-shared", "-o", LIB] + objs + ["-gencode", "arch=compute_100a,code=sm_100a", "-ldl"])
    return LIB


def build_host(verbose=False):
    srcs = [os.path.join(HOST, f) for f in sorted(os.listdir(HOST)) if f.endswith(".cpp")] if os.path.isdir(HOST) else []
    if not srcs:
        return None
    deps = srcs + [os.path.join(HOST, f) for f in os.listdir(HOST) if f.endswith(".hpp")]
    deps.append(os.path.join(ROOT, "include", "jubann.h"))
    if _newer(HOST_LIB, deps):
        cxx = _cxx()
        _run([cxx, "-O2", "-std=c++17", "-fPIC", "-shared", "-pthread", "-Wall", "-I", os.path.join(ROOT, "include"),
              "-o", HOST_LIB] + srcs + ["-ldl"])
    return HOST_LIB


def build_oracle():
    """Test infrastructure: the CPU oracle (building the checker is not using it)."""
    _run(["make", "-C", os.path.join(ROOT, "oracle"), "CXX=" + _cxx()])
    # oracle/_ref: the reference's own crc32.cpp, compiled where it lies (only in the container
    # that has /root/reference; the GPU box uses the prebuilt file that travelled with the snapshot)
    _run(["make", "-C", os.path.join(ROOT, "oracle"), "CXX=" + _cxx(), "ref"])


def build_all(verbose=False):
    lib = build_cuda(verbose)
    build_host(verbose)
    build_oracle()
    return lib


if __name__ == "__main__":
    print(build_all(verbose="-v" in sys.argv))
