"""Loader / builder of libmm25.so — the C-ABI library declared in include/mm25.h.

The product path has no CPU fallback: if the shared library is missing or the device is not a
B200 (sm_100), calls raise.  `build()` cross-compiles with nvcc (no GPU needed).
"""
from __future__ import annotations

import ctypes
import os
import subprocess
import threading

_HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(_HERE, "csrc")
LIB_DIR = os.path.join(_HERE, "lib")
LIB_PATH = os.path.join(LIB_DIR, "libmm25.so")
INCLUDE = os.path.join(os.path.dirname(_HERE), "include")

SOURCES = ["prep.cu", "sim_simt.cu", "sim_tc.cu", "sim_tc_bwd.cu", "sim_tc_mod.cu", "gemm_tc.cu", "gemm_tc2.cu",
           "gemm_te.cu",
           "api_sim.cu", "hopfield.cu", "samplers.cu", "beta_vmf.cu", "bessel.cu", "topk.cu", "rank.cu", "sim_tc_rank.cu"]

_lock = threading.Lock()
_lib = None


class MM25Error(RuntimeError):
    pass


def _sources():
    return [os.path.join(CSRC, s) for s in SOURCES if os.path.exists(os.path.join(CSRC, s))]


def _stale() -> bool:
    if not os.path.exists(LIB_PATH):
        return True
    t = os.path.getmtime(LIB_PATH)
    deps = _sources() + [os.path.join(CSRC, f) for f in os.listdir(CSRC)
                         if f.endswith((".cuh", ".h", ".inc"))] + [os.path.join(INCLUDE, "mm25.h")]
    return any(os.path.getmtime(p) > t for p in deps)


def build(force: bool = False, verbose: bool = False) -> str:
    """Compile every .cu under csrc/ into lib/libmm25.so for sm_100a."""
    with _lock:
        if not force and not _stale():
            return LIB_PATH
        os.makedirs(LIB_DIR, exist_ok=True)
        nvcc = os.environ.get("NVCC", "nvcc")
        objs = []
        procs = []
        obj_dir = os.path.join(LIB_DIR, "obj")
        os.makedirs(obj_dir, exist_ok=True)
        for src in _sources():
            obj = os.path.join(obj_dir, os.path.basename(src) + ".o")
            objs.append(obj)
            cmd = [nvcc, "-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3",
                   "-std=c++17", "-Xcompiler", "-fPIC", "-I", INCLUDE, "-c", src, "-o", obj]
            if verbose:
                cmd.insert(1, "-Xptxas")
                cmd.insert(2, "-v")
            procs.append((cmd, subprocess.Popen(cmd, stdout=subprocess.PIPE,
                                                stderr=subprocess.STDOUT, text=True)))
        for cmd, p in procs:
            out, _ = p.communicate()
            if p.returncode != 0:
                raise MM25Error("nvcc failed: %s\n%s" % (" ".join(cmd), out))
            if verbose and out:
                print(out)
        cmd = [nvcc, "-shared", "-o", LIB_PATH + ".tmp"] + objs
        r = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
        if r.returncode != 0:
            raise MM25Error("link failed: %s\n%s" % (" ".join(cmd), r.stdout))
        os.replace(LIB_PATH + ".tmp", LIB_PATH)
        return LIB_PATH


_i64, _int, _f32, _vp, _sz = (ctypes.c_int64, ctypes.c_int, ctypes.c_float, ctypes.c_void_p,
                              ctypes.c_size_t)
_u64, _f64 = ctypes.c_uint64, ctypes.c_double

# name -> (restype, argtypes); must list every symbol include/mm25.h declares
SIGNATURES = {
    "mm25_version": (ctypes.c_char_p, []),
    "mm25_last_cuda_error": (ctypes.c_char_p, []),
    "mm25_device_ok": (_int, []),
    "mm25_launch_count": (ctypes.c_longlong, []),
    "mm25_debug_timing": (None, [_int]),
    "mm25_debug_timing_read": (_int, [_int, _vp, _vp]),
    "mm25_debug_force_slow_epilogue": (None, [_int]),
    "mm25_debug_set_g_panel_bytes": (None, [_sz]),
    "mm25_debug_gemm_bf16": (_int, [_vp, _int, _vp, _int, _vp, _i64, _i64, _i64, _f32, _int, _vp]),
    "mm25_debug_gemm_grouped": (_int, [_vp, _int, _vp, _int, _vp, _i64, _i64, _vp, _int, _vp, _int,
                                       _vp, _i64, _i64, _i64, _vp]),
    "mm25_debug_gemm_scatter": (_int, [_vp, _int, _vp, _int, _vp, _int, _i64, _int, _i64, _i64, _i64,
                                       _vp]),
    "mm25_debug_gemm_te": (_int, [_vp, _i64, _i64, _vp, _vp, _vp, _vp, _vp, _i64, _vp, _vp, _vp,
                                  _vp, _i64, _i64, _vp]),
    "mm25_debug_te_stats": (_int, [_vp, _int]),
    "mm25_debug_gemm_schedule": (_int, [_i64, _i64, _i64, _i64, _int, _vp]),
    "mm25_rownorm_colsum_floats": (_i64, [_i64, _i64, _i64]),
    "mm25_rownorm_fwd": (_int, [_vp, _int, _vp, _int, _vp, _vp, _i64, _i64, _i64, _int, _f32, _vp]),
    "mm25_rownorm_colsum_noalloc_floats": (_i64, [_i64, _i64, _i64]),
    "mm25_rownorm_colsum_noalloc": (_int, [_vp, _int, _vp, _int, _vp, _vp, _i64, _i64, _i64, _int,
                                           _f32, _vp]),
    "mm25_rownorm_bwd": (_int, [_vp, _vp, _int, _vp, _vp, _int, _i64, _i64, _i64, _int, _vp, _int,
                                _vp]),
    "mm25_sim_fwd_workspace_bytes": (_sz, [_i64, _i64, _i64, _i64, _int, _int]),
    "mm25_sim_fwd": (_int, [_vp, _vp, _int, _vp, _vp, _vp, _i64, _i64, _i64, _i64, _i64, _vp, _int,
                            _f32, _int, _i64, _vp, _vp, _vp, _vp, _vp, _sz, _int, _f32, _vp, _vp,
                            _vp]),
    "mm25_sim_bwd_workspace_bytes": (_sz, [_i64, _i64, _i64, _i64, _int, _int, _int]),
    "mm25_sim_bwd_nparts": (_i64, [_i64, _i64, _i64, _i64, _int, _int]),
    "mm25_debug_last_sim_path": (_int, []),
    "mm25_sim_mod_workspace_bytes": (_sz, [_i64, _i64, _i64, _i64, _i64, _int, _int]),
    "mm25_sim_mod_bwd_nparts": (_i64, [_i64, _i64, _i64, _i64, _i64, _int]),
    "mm25_sim_bwd": (_int, [_vp, _vp, _int, _vp, _vp, _vp, _i64, _i64, _i64, _i64, _i64, _vp, _int,
                            _f32, _int, _i64, _vp, _vp, _f32, _f32, _f32, _f32, _vp, _vp, _vp, _vp,
                            _vp, _vp, _sz, _int, _f32, _vp, _vp, _vp]),
    "mm25_sample_projnormal_fwd": (_int, [_vp, _vp, _vp, _u64, _i64, _i64, _i64, _vp, _vp, _vp]),
    "mm25_sample_projnormal_bwd": (_int, [_vp, _vp, _vp, _vp, _u64, _i64, _i64, _i64, _vp, _vp,
                                          _vp]),
    "mm25_sphere_sample_fwd": (_int, [_vp, _vp, _vp, _u64, _i64, _i64, _i64, _f32, _f32, _f32,
                                      _f32, _f32, _vp, _vp]),
    "mm25_sphere_sample_bwd": (_int, [_vp, _vp, _vp, _vp, _i64, _i64, _i64, _f32, _f32, _f32,
                                      _f32, _f32, _vp, _vp, _vp]),
    "mm25_vmf_sample_w": (_int, [_vp, _i64, _i64, _i64, _u64, _vp, _vp, _i64, _vp, _vp, _vp]),
    "mm25_beta_sample": (_int, [_vp, _vp, _i64, _i64, _u64, _vp, _vp]),
    "mm25_beta_grad": (_int, [_vp, _vp, _vp, _i64, _i64, _vp, _vp]),
    "mm25_log_bessel": (_int, [_vp, _i64, _f64, _vp, _vp, _vp]),
    "mm25_sim_bwd_rs": (_int, [_vp, _vp, _int, _i64, _i64, _i64, _vp, _int, _f32, _i64, _vp, _vp,
                               _f32, _f32, _f32, _f32, _vp, _vp, _vp, _i64, _vp, _vp, _sz, _f32,
                               _vp, _vp, _vp]),
    "mm25_topk_rows": (_int, [_vp, _i64, _i64, _i64, _i64, _vp, _vp, _vp]),
    "mm25_topk_workspace_bytes": (_sz, [_i64, _i64, _i64]),
    "mm25_topk_rows_ws": (_int, [_vp, _i64, _i64, _i64, _i64, _vp, _vp, _vp, _sz, _vp]),
    "mm25_rank_both": (_int, [_vp, _i64, _i64, _i64, _i64, _vp, _vp, _vp, _vp, _vp, _vp, _vp]),
    "mm25_split_width": (_int, []),
    "mm25_split_bf16x2": (_int, [_vp, _i64, _i64, _int, _vp, _vp, _vp, _vp]),
    "mm25_sim_split_eligible": (_int, [_i64, _i64, _i64, _i64]),
    "mm25_sim_bwd_split_workspace_bytes": (_sz, [_i64, _i64]),
    "mm25_sim_bwd_split_nparts": (_i64, [_i64, _i64, _i64]),
    "mm25_sim_bwd_split": (_int, [_vp, _vp, _vp, _vp, _vp, _vp, _i64, _i64, _i64, _i64, _vp, _int,
                                  _f32, _i64, _vp, _vp, _f32, _f32, _f32, _f32, _vp, _vp, _vp, _vp,
                                  _vp, _sz, _f32, _int, _vp]),
    "mm25_hopfield_split_eligible": (_int, [_i64, _i64, _i64]),
    "mm25_hopfield_bwd_split_workspace_bytes": (_sz, [_i64, _i64, _i64]),
    "mm25_hopfield_bwd_split": (_int, [_vp, _vp, _vp, _vp, _vp, _vp, _vp, _i64, _i64, _i64, _vp, _vp,
                                       _vp, _vp, _vp, _vp, _sz, _vp]),
    "mm25_sim_rank": (_int, [_vp, _vp, _int, _i64, _i64, _i64, _vp, _vp, _vp, _vp]),
    "mm25_rowdot": (_int, [_vp, _vp, _vp, _i64, _i64, _vp]),
    "mm25_hopfield_bwd_pass": (_int, [_vp, _vp, _vp, _vp, _vp, _vp, _i64, _i64, _i64, _f32, _vp,
                                      _vp, _int, _vp, _vp]),
    "mm25_hopfield_bwd_tc_workspace_bytes": (_sz, [_i64, _i64, _i64]),
    "mm25_hopfield_bwd_tc": (_int, [_vp, _vp, _vp, _i64, _i64, _i64, _vp, _vp, _vp, _vp, _vp, _vp,
                                    _sz, _vp]),
    "mm25_sum_parts": (_int, [_vp, _i64, _f32, _vp, _vp]),
    "mm25_col_merge": (_int, [_vp, _vp, _i64, _i64, _vp, _vp]),
    "mm25_col_merge_ld": (_int, [_vp, _vp, _i64, _i64, _i64, _vp, _vp]),
    "mm25_peer_bcast": (_int, [_vp, _i64, _vp, _i64, _i64, _vp]),
    "mm25_peer_bcast_mc": (_int, [_vp, _i64, _vp, _i64, _vp]),
    "mm25_smooth_term": (_int, [_vp, _i64, _i64, _i64, _i64, _i64, _vp, _int, _f32, _f32, _vp,
                                _vp]),
    "mm25_sharded_loss": (_int, [_vp, _i64, _i64, _i64, _vp, _i64, _i64, _i64, _i64, _vp, _int,
                                 _f32, _f32, _vp, _vp]),
    "mm25_scale_param_grad": (_int, [_vp, _vp, _int, _f32, _vp, _vp]),
    "mm25_clip_loss_finalize": (_int, [_vp, _vp, _vp, _vp, _vp, _i64, _i64, _i64, _i64, _i64, _vp,
                                       _int, _f32, _f32, _vp, _vp]),
}

_ERRORS = {-1: "bad/unsupported shape", -2: "unsupported dtype", -3: "misaligned pointer",
           -4: "workspace too small", -5: "device is not sm_100", -6: "CUDA error",
           -7: "bad argument"}


def lib() -> ctypes.CDLL:
    """Return the loaded library; raises MM25Error when it is missing (no fallback)."""
    global _lib
    if _lib is not None:
        return _lib
    with _lock:
        if _lib is not None:
            return _lib
        if not os.path.exists(LIB_PATH):
            raise MM25Error(
                "libmm25.so not built (%s). Run `python -c 'import __graft_entry__ as g; "
                "g.build()'`. There is no CPU / PyTorch fallback for this path." % LIB_PATH)
        handle = ctypes.CDLL(LIB_PATH)
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(handle, name)
            fn.restype = res
            fn.argtypes = args
        _lib = handle
    return _lib


def check(rc: int, what: str) -> None:
    if rc != 0:
        msg = _ERRORS.get(rc, "error %d" % rc)
        if rc == -6:
            msg += ": " + lib().mm25_last_cuda_error().decode()
        raise MM25Error("%s failed: %s" % (what, msg))
