"""The C++ binding INTEGRATION.md shows must keep compiling against the host header."""
import os
import shutil
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_reference_side_binding_compiles():
    cxx = shutil.which("g++")
    if cxx is None:
        pytest.skip("no g++ on this machine")
    cmd = [cxx, "-std=c++17", "-fsyntax-only", "-Wall", "-I", os.path.join(ROOT, "include"),
           "-I", os.path.join(ROOT, "jubatus_b200", "host"), os.path.join(ROOT, "tests", "cpp", "integration_snippet.cpp")]
    r = subprocess.run(cmd, capture_output=True, text=True)
    assert r.returncode == 0, r.stderr[-4000:]
