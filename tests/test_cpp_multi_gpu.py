"""Builds and runs tests/cpp/multi_gpu_test.cpp: a C++ process (no Python inside) that shards one
table over every GPU of the box through jubann_create_multi and the server facade, bit-exact against
the one-GPU handle and a brute-force ranking.  On a one-GPU box it runs with a group of one device."""
import os
import shutil
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.gpu
def test_cpp_host_shards_over_the_gpus_of_the_box(tmp_path):
    import torch
    cxx = shutil.which("g++")
    if cxx is None:
        pytest.skip("no g++ on this machine")
    exe = str(tmp_path / "multi_gpu_test")
    cmd = [cxx, "-std=c++17", "-O2", "-Wall", "-I", os.path.join(ROOT, "include"), "-I", os.path.join(ROOT, "jubatus_b200", "host"),
           os.path.join(ROOT, "tests", "cpp", "multi_gpu_test.cpp"), "-o", exe, "-ldl", "-pthread"]
    r = subprocess.run(cmd, capture_output=True, text=True)
    assert r.returncode == 0, r.stderr[-4000:]
    ndev = min(8, torch.cuda.device_count())
    r = subprocess.run([exe, os.path.join(ROOT, "jubatus_b200", "libjubann.so"), str(ndev)], capture_output=True, text=True,
                       timeout=600)
    assert r.returncode == 0, (r.stdout + r.stderr)[-4000:]
    assert r.stdout.strip().endswith("ok")
