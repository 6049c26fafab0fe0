"""ctypes binding of the product library ``libminiopt_b200.so`` (CUDA kernels + C ABI, include/mini_opt_b200.h).

There is no CPU fallback: every compute call goes to the CUDA path and raises if the library is missing or the
device is not a B200-class (sm_100) GPU. Nothing under ``oracle/`` is ever imported from here.
"""
from __future__ import annotations

import ctypes as C
import subprocess
from pathlib import Path

from . import abi

import os

PKG = Path(__file__).resolve().parent
# MINI_OPT_B200_LIB: harness-only override used by the profiling scripts under tools/ to load an instrumented build
# (e.g. -DMO_WD_TRACE); the product C/C++ callers link libminiopt_b200.so directly.
LIB_PATH = Path(os.environ.get("MINI_OPT_B200_LIB", PKG / "libminiopt_b200.so"))

EXPORTS = [
    "mo_qp_params_default", "mo_qp_vars_stride", "mo_qp_ip_solve_batch", "mo_qp_ip_solve_batch_host", "mo_qp_ip_solve_batch_host_multi",
    "mo_qp_ip_newton_step_batch", "mo_qp_nullspace_solve_batch", "mo_qp_nullspace_solve_batch_noalloc", "mo_qp_nullspace_solve_batch_host", "mo_qp_kernel_variant", "mo_kernel_launch_count", "mo_last_error_string",
    "mo_abi_version", "mo_host_alloc", "mo_host_alloc_interleaved", "mo_host_free", "mo_measure_fp64_peak", "mo_measure_dmma_peak", "mo_nls_family_dims", "mo_nls_solve_family_batch", "mo_debug_set_tma", "mo_debug_set_tpp_async", "mo_debug_set_dm_kernel", "mo_debug_set_wd_lock", "mo_debug_set_ik_warps", "mo_debug_set_ik_sync",
    "mo_nls_params_default", "mo_nls_linearize_batch", "mo_chain_fk_batch", "mo_nls_solve_chain_ik_batch",
]


class MiniOptError(RuntimeError):
    def __init__(self, code: int, msg: str):
        super().__init__(f"mini_opt_b200 error {code}: {msg}")
        self.code = code


def build(verbose: bool = False) -> None:
    """Compile the CUDA extension in-tree for sm_100a (nvcc cross-compiles without a GPU)."""
    import os
    jobs = str(max(1, min(5, os.cpu_count() or 1)))  # five translation units; the three big ones take ~5 minutes each
    res = subprocess.run(["make", "-j", jobs, "-C", str(PKG / "csrc"), "all"], capture_output=True, text=True)
    if verbose or res.returncode != 0:
        print(res.stdout[-4000:], res.stderr[-4000:])
    if res.returncode != 0:
        raise RuntimeError("building libminiopt_b200.so failed")


_lib = None


def load():
    """Load the product library; raises (loudly) if it has not been built."""
    global _lib
    if _lib is not None:
        return _lib
    if not LIB_PATH.exists():
        raise FileNotFoundError(
            f"{LIB_PATH} is missing: run `python -c 'import __graft_entry__ as g; g.build()'` "
            "(there is no CPU fallback for the solve path)")
    lib = C.CDLL(str(LIB_PATH))
    BP, PP, OP = C.POINTER(abi.MoQpBatch), C.POINTER(abi.MoQpParams), C.POINTER(abi.MoQpOut)
    lib.mo_qp_params_default.argtypes = [PP]
    lib.mo_qp_params_default.restype = None
    lib.mo_qp_vars_stride.argtypes = [C.c_int32] * 3
    lib.mo_qp_vars_stride.restype = C.c_int64
    lib.mo_qp_ip_solve_batch.argtypes = [BP, PP, OP, C.c_int, C.c_void_p]
    lib.mo_qp_ip_solve_batch.restype = C.c_int
    lib.mo_qp_ip_solve_batch_host.argtypes = [BP, PP, OP, C.c_int]
    lib.mo_qp_ip_solve_batch_host.restype = C.c_int
    lib.mo_qp_ip_solve_batch_host_multi.argtypes = [BP, PP, OP, C.POINTER(C.c_int), C.c_int]
    lib.mo_qp_ip_solve_batch_host_multi.restype = C.c_int
    lib.mo_qp_ip_newton_step_batch.argtypes = [BP, C.c_double, C.c_int32] + [C.c_void_p] * 5 + [C.c_int, C.c_void_p]
    lib.mo_qp_ip_newton_step_batch.restype = C.c_int
    lib.mo_qp_nullspace_solve_batch.argtypes = [BP, C.c_void_p, C.c_void_p, C.c_int, C.c_void_p]
    lib.mo_qp_nullspace_solve_batch.restype = C.c_int
    lib.mo_qp_nullspace_solve_batch_noalloc.argtypes = [BP, C.c_void_p, C.c_void_p, C.c_int, C.c_void_p]
    lib.mo_qp_nullspace_solve_batch_noalloc.restype = C.c_int
    lib.mo_qp_nullspace_solve_batch_host.argtypes = [BP, C.c_void_p, C.c_void_p, C.c_int]
    lib.mo_qp_nullspace_solve_batch_host.restype = C.c_int
    lib.mo_qp_kernel_variant.argtypes = [C.c_int32] * 4
    lib.mo_qp_kernel_variant.restype = C.c_char_p
    lib.mo_kernel_launch_count.restype = C.c_int64
    lib.mo_last_error_string.restype = C.c_char_p
    lib.mo_abi_version.restype = C.c_int
    lib.mo_host_alloc.argtypes = [C.c_size_t]
    lib.mo_host_alloc.restype = C.c_void_p
    lib.mo_host_free.argtypes = [C.c_void_p]
    lib.mo_host_free.restype = None
    lib.mo_measure_fp64_peak.argtypes = [C.c_int, C.POINTER(C.c_double)]
    lib.mo_measure_fp64_peak.restype = C.c_int
    lib.mo_debug_set_ik_sync.argtypes = [C.c_int]
    lib.mo_debug_set_ik_sync.restype = C.c_int
    lib.mo_debug_set_ik_warps.argtypes = [C.c_int]
    lib.mo_debug_set_ik_warps.restype = C.c_int
    lib.mo_debug_set_dm_kernel.argtypes = [C.c_int]
    lib.mo_debug_set_dm_kernel.restype = C.c_int
    lib.mo_debug_set_tpp_async.argtypes = [C.c_int]
    lib.mo_debug_set_tpp_async.restype = C.c_int
    lib.mo_debug_set_tma.argtypes = [C.c_int]
    lib.mo_debug_set_tma.restype = C.c_int
    lib.mo_measure_dmma_peak.argtypes = [C.c_int, C.POINTER(C.c_double)]
    lib.mo_measure_dmma_peak.restype = C.c_int
    lib.mo_nls_params_default.argtypes = [C.POINTER(abi.MoNlsParams)]
    lib.mo_nls_params_default.restype = None
    lib.mo_nls_linearize_batch.argtypes = [C.POINTER(abi.MoNlsLinearizeArgs), C.c_int, C.c_void_p]
    lib.mo_nls_linearize_batch.restype = C.c_int
    lib.mo_chain_fk_batch.argtypes = [C.POINTER(abi.MoChainIkBatch)] + [C.c_void_p] * 5 + [C.c_int, C.c_void_p]
    lib.mo_chain_fk_batch.restype = C.c_int
    lib.mo_nls_solve_chain_ik_batch.argtypes = [C.POINTER(abi.MoChainIkBatch), C.POINTER(abi.MoNlsParams),
                                                C.POINTER(abi.MoNlsOut), C.c_int, C.c_void_p]
    lib.mo_nls_solve_chain_ik_batch.restype = C.c_int
    lib.mo_host_alloc_interleaved.argtypes = [C.c_size_t]
    lib.mo_host_alloc_interleaved.restype = C.c_void_p
    lib.mo_nls_family_dims.argtypes = [C.c_int32] + [C.POINTER(C.c_int32)] * 3
    lib.mo_nls_family_dims.restype = C.c_int
    lib.mo_nls_solve_family_batch.argtypes = [C.POINTER(abi.MoNlsFamilyBatch), C.POINTER(abi.MoNlsParams),
                                              C.POINTER(abi.MoNlsOut), C.c_int, C.c_void_p]
    lib.mo_nls_solve_family_batch.restype = C.c_int
    if lib.mo_abi_version() != abi.MO_ABI_VERSION:
        raise RuntimeError("libminiopt_b200.so ABI version mismatch; rebuild")
    _lib = lib
    return lib


def check(rc: int) -> None:
    if rc != 0:
        msg = load().mo_last_error_string().decode() if rc in (abi.MO_ERR_CUDA, abi.MO_ERR_NO_DEVICE,
                                                               abi.MO_ERR_UNSUPPORTED_SHAPE) else ""
        names = {-1: "invalid argument", -2: "invalid params", -3: "unsupported shape", -4: "CUDA error",
                 -5: "no sm_100 device"}
        raise MiniOptError(rc, f"{names.get(rc, '?')} {msg}".strip())


def kernel_variant(n: int, k: int, m: int, flags: int) -> str:
    return load().mo_qp_kernel_variant(n, k, m, flags).decode()


def launch_count() -> int:
    return int(load().mo_kernel_launch_count())


def measure_dmma_peak(device_index: int = 0) -> float:
    """FP64 tensor-path (DMMA.8x8x4) peak of this GPU in TFLOP/s (live microbenchmark)."""
    v = C.c_double(0.0)
    check(load().mo_measure_dmma_peak(int(device_index), C.byref(v)))
    return float(v.value)


def measure_fp64_peak(device_index: int = 0) -> float:
    """FP64 DFMA peak of this GPU in TFLOP/s (live microbenchmark)."""
    v = C.c_double(0.0)
    check(load().mo_measure_fp64_peak(int(device_index), C.byref(v)))
    return float(v.value)
