"""ctypes loader for libweclip_sm100.so (the C ABI declared in include/weclip_b200.h).

The product path has no CPU or PyTorch fallback: a missing library or a non-sm_100 device raises.
"""
from __future__ import annotations

import ctypes
import os
import re
import subprocess
from ctypes import c_char_p, c_float, c_int, c_int64, c_size_t, c_void_p, POINTER

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libweclip_sm100.so")
HEADER_PATH = os.path.join(os.path.dirname(_HERE), "include", "weclip_b200.h")

_lib = None


class WeclipError(RuntimeError):
    pass


def build(verbose: bool = False) -> str:
    """Compile every .cu under csrc/ for sm_100a into libweclip_sm100.so (nvcc cross-compiles without a GPU)."""
    cmd = ["make", "-C", os.path.join(_HERE, "csrc"), "-j", str(os.cpu_count() or 4)]
    res = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
    if verbose or res.returncode != 0:
        print(res.stdout)
    if res.returncode != 0:
        raise WeclipError("building libweclip_sm100.so failed")
    return LIB_PATH


def declared_symbols() -> list:
    """Names of every function include/weclip_b200.h declares (used by the export test)."""
    src = open(HEADER_PATH).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(wc_[a-z0-9_]+)\s*\(", src)))


def lib() -> ctypes.CDLL:
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise WeclipError(
                f"{LIB_PATH} is missing: run `python -c 'import __graft_entry__ as g; g.build()'` "
                "(there is no CPU / PyTorch fallback for the hot path)")
        _lib = ctypes.CDLL(LIB_PATH)
        _lib.wc_last_error.restype = c_char_p
        for name in declared_symbols():
            if name.endswith("_workspace_bytes") and hasattr(_lib, name):
                getattr(_lib, name).restype = c_size_t
    return _lib


def check(rc: int, what: str = ""):
    if rc != 0:
        raise WeclipError(f"{what} failed ({rc}): {lib().wc_last_error().decode()}")


def require_device():
    """Fail loudly when the CUDA path cannot run (no fallback)."""
    import torch
    if not torch.cuda.is_available():
        raise WeclipError("weclip_b200 needs a CUDA device (sm_100a); no CPU fallback exists")
    if not lib().wc_device_ok():
        raise WeclipError("weclip_b200 kernels are built for sm_100a (B200) only")


def ptr(t) -> c_void_p:
    return c_void_p(t.data_ptr()) if t is not None else c_void_p(0)


def stream_ptr() -> c_void_p:
    """torch's current stream as a raw cudaStream_t.  Called once per kernel launch: torch.cuda.current_stream() builds a Stream
    object (device-index lookups, ~14 us measured); the raw getter is two orders of magnitude cheaper."""
    import torch
    raw = getattr(torch._C, "_cuda_getCurrentRawStream", None)
    if raw is not None:
        return c_void_p(raw(torch.cuda.current_device()))
    return c_void_p(torch.cuda.current_stream().cuda_stream)


def int_array(vals):
    return (c_int * len(vals))(*[int(v) for v in vals])
