"""CPU (nvcc cross-compile, no GPU): build-quality gates for the hot tensor-core kernels — no
register spills, the register budgets the launch bounds assume, and the tcgen05 / TMA SASS
mnemonics that prove the path (B200_PROFILING.md) are present in the shipped objects."""
import os
import re
import shutil
import subprocess
from concurrent.futures import ThreadPoolExecutor

import pytest

from conftest import ROOT

CSRC = os.path.join(ROOT, "multimodal-2025_b200", "csrc")
HOT = {  # file -> kernels that must not spill
    "sim_tc.cu": ["sim_fwd_tc_kernel"],
    "sim_tc_bwd.cu": ["sim_grad_tc_kernel"],
    "sim_tc_mod.cu": ["sim_mod_tc_kernel"],
    "gemm_tc.cu": ["gemm_grouped_kernel"],
    "rank.cu": ["rank_both_kernel"],
}


def _ptxas_v(src):
    cmd = ["nvcc", "-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-std=c++17", "-Xptxas",
           "-v", "-I", os.path.join(ROOT, "include"), "-c", os.path.join(CSRC, src), "-o",
           os.devnull]
    r = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
    assert r.returncode == 0, r.stdout[-2000:]
    return r.stdout


@pytest.mark.skipif(shutil.which("nvcc") is None, reason="no nvcc")
def test_hot_kernels_do_not_spill():
    with ThreadPoolExecutor(len(HOT)) as ex:
        outs = dict(zip(HOT, ex.map(_ptxas_v, HOT)))
    for src, kernels in HOT.items():
        blocks = re.split(r"Compiling entry function '", outs[src])[1:]
        seen = set()
        for b in blocks:
            name = b.split("'")[0]
            for k in kernels:
                if k in name:
                    seen.add(k)
                    m = re.search(r"(\d+) bytes spill stores, (\d+) bytes spill loads", b)
                    assert m and m.group(1) == "0" and m.group(2) == "0", (src, name, m.group(0))
                    regs = int(re.search(r"Used (\d+) registers", b).group(1))
                    assert regs <= 255
        assert seen == set(kernels), (src, seen)


@pytest.mark.skipif(shutil.which("cuobjdump") is None, reason="no cuobjdump")
def test_shipped_objects_contain_tcgen05_and_tma_sass(lib_built):
    """UTCHMMA (tcgen05.mma), UTMALDG (TMA load), LDTM (tcgen05.ld) in the sweep and GEMM objects
    of the library that ships to the GPU box."""
    obj_dir = os.path.join(ROOT, "multimodal-2025_b200", "lib", "obj")
    for src in ("sim_tc.cu", "sim_tc_bwd.cu", "gemm_tc.cu"):
        obj = os.path.join(obj_dir, src + ".o")
        assert os.path.exists(obj), obj
        sass = subprocess.run(["cuobjdump", "-sass", obj], stdout=subprocess.PIPE, text=True).stdout
        for mnemonic in ("UTCHMMA", "UTMALDG", "LDTM"):
            assert mnemonic in sass, (src, mnemonic)
