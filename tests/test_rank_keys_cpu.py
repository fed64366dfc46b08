"""CPU: the ordering helpers the rank kernel compiles (csrc/rank_keys.h), built for the host
with g++ and checked against an independently written stable descending sort on rows full of
ties, signed zeros, infinities and NaNs (tests/rank_keys_host_test.cc)."""
import os
import shutil
import subprocess

import pytest

from conftest import ROOT


def test_rank_keys_host_build_matches_stable_sort(tmp_path):
    gxx = shutil.which("g++")
    if gxx is None:
        pytest.skip("no g++")
    exe = str(tmp_path / "rank_keys_test")
    src = os.path.join(ROOT, "tests", "rank_keys_host_test.cc")
    r = subprocess.run([gxx, "-O2", "-std=c++17", "-o", exe, src], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    r = subprocess.run([exe], capture_output=True, text=True, timeout=120)
    assert r.returncode == 0 and "rank keys ok" in r.stdout, r.stdout + r.stderr
