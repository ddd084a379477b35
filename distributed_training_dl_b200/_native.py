"""ctypes binding of ``libdtdl_b200.so`` (the sm_100a kernels + peer-memory runtime).

The prototypes are read from ``include/dtdl_b200.h`` so Python and C cannot drift
apart.  The library is built in-tree by ``build.sh`` / ``__graft_entry__.build()``.
On a machine with a GPU a missing library is a hard error: the CUDA path must be
the one that runs, never a silent PyTorch fallback.
"""
from __future__ import annotations

import ctypes
import os
import re
import subprocess
import threading

_HERE = os.path.dirname(os.path.abspath(__file__))
_ROOT = os.path.dirname(_HERE)
# DTDL_LIB: load another build of the library (A/B measurements of kernel variants)
LIB_PATH = os.environ.get("DTDL_LIB") or os.path.join(_HERE, "lib", "libdtdl_b200.so")
HEADER_PATH = os.path.join(_ROOT, "include", "dtdl_b200.h")

_lock = threading.Lock()
_lib = None


class NativeError(RuntimeError):
    pass


_CTYPE = {
    "int": ctypes.c_int,
    "long long": ctypes.c_longlong,
    "float": ctypes.c_float,
}


def _parse_header(path):
    """Returns {name: (restype, [argtypes])} for every prototype in the header."""
    src = open(path).read()
    src = re.sub(r"/\*.*?\*/", " ", src, flags=re.S)
    protos = {}
    for m in re.finditer(r"\b(const char\*|int|long long)\s+(dtdl_\w+)\s*\(([^;]*?)\)\s*;", src, flags=re.S):
        ret, name, args = m.group(1), m.group(2), m.group(3)
        restype = ctypes.c_char_p if ret == "const char*" else _CTYPE[ret]
        argtypes = []
        args = " ".join(args.split())
        if args and args != "void":
            for a in args.split(","):
                a = a.strip()
                if "*" in a:
                    argtypes.append(ctypes.c_void_p)
                else:
                    base = "long long" if a.startswith("long long") else a.split()[0]
                    argtypes.append(_CTYPE[base])
        protos[name] = (restype, argtypes)
    return protos


def build(verbose=False):
    """Compile the library in-tree (nvcc cross-compiles for sm_100a without a GPU)."""
    script = os.path.join(_ROOT, "build.sh")
    res = subprocess.run(["bash", script], cwd=_ROOT, capture_output=True, text=True)
    if verbose or res.returncode != 0:
        print(res.stdout)
        print(res.stderr)
    if res.returncode != 0:
        raise NativeError("building libdtdl_b200.so failed:\n" + res.stdout + res.stderr)
    return LIB_PATH


def available():
    return os.path.exists(LIB_PATH)


def lib():
    """Loads (once) and returns the ctypes library handle with typed prototypes."""
    global _lib
    if _lib is not None:
        return _lib
    with _lock:
        if _lib is not None:
            return _lib
        if not os.path.exists(LIB_PATH):
            raise NativeError(
                f"{LIB_PATH} is missing; run ./build.sh (or __graft_entry__.build()) first. "
                "The framework does not fall back to PyTorch kernels on a GPU.")
        handle = ctypes.CDLL(LIB_PATH, mode=ctypes.RTLD_GLOBAL)
        for name, (restype, argtypes) in _parse_header(HEADER_PATH).items():
            fn = getattr(handle, name)
            fn.restype = restype
            fn.argtypes = argtypes
        _lib = handle
    return _lib


def check(rc):
    if rc != 0:
        msg = lib().dtdl_last_error()
        raise NativeError(msg.decode() if msg else f"native call failed with code {rc}")


# kernels launched per native call (everything not listed launches none)
_LAUNCHES = {"conv_fprop": 1, "conv_fprop_stats": 1, "conv_fprop_gather": 1, "conv_dgrad_s2": 4, "comm_nvls_allreduce": 1, "comm_reduce_update16": 1, "comm_nvls_reduce_update16": 1, "cast_f32_bf16": 1,
             "comm_nvls_reduce_update": 1, "comm_nvls_broadcast": 1, "conv_wgrad": 2, "pack_weights": 1,
             "nchw_to_nhwc": 1, "nchw_to_s2d": 1, "bn_stats": 1, "bn_apply": 1, "bn_apply_stats": 1, "bn_bwd": 2,
             "bn_bwd_add": 2, "bn_bwd_fused": 1, "relu_bwd": 1, "add": 1, "add_prefix": 1, "pool_fwd": 1, "pool_bwd": 1,
             "subsample2": 1, "gap_fwd": 1, "gap_bwd": 1, "softmax_ce": 1, "colsum": 1, "pad_cast": 1,
             "unpad_cast": 1, "linear_f32_fwd": 1, "linear_f32_bwd": 2, "comm_barrier": 1, "comm_allreduce": 1,
             "comm_reduce_update": 1, "comm_broadcast": 1, "comm_allgather": 1, "local_update": 1}
_launch_count = 0


def reset_launch_count():
    global _launch_count
    _launch_count = 0


def add_launches(n):
    """Kernels launched by a native call beyond its ``_LAUNCHES`` entry (the one-CTA flag-exchange
    kernels around a collective, which depend on world size and mode)."""
    global _launch_count
    _launch_count += n


def launch_count():
    """Number of this library's kernels launched since ``reset_launch_count()``."""
    return _launch_count


def call(name, *args):
    """Calls ``dtdl_<name>`` and raises NativeError on a non-zero return code."""
    global _launch_count
    fn = getattr(lib(), "dtdl_" + name)
    check(fn(*args))
    _launch_count += _LAUNCHES.get(name, 0)


def ptr(t):
    """Device pointer of a tensor (or 0 for None) as an int usable for c_void_p."""
    return 0 if t is None else t.data_ptr()


def stream_ptr():
    import torch
    return torch.cuda.current_stream().cuda_stream
