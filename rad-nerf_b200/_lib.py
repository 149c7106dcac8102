"""ctypes binding of libradnerf_b200.so (C ABI declared in include/radnerf_b200.h).

The library is plain CUDA (no torch types cross the boundary): this module hands it raw
device pointers (`tensor.data_ptr()`), sizes and the current CUDA stream.  There is NO CPU
fallback: if the shared library is missing, every operator raises ``RuntimeError``.
"""
from __future__ import annotations

import ctypes
import os
import re

import torch

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "lib", "libradnerf_b200.so")
HEADER_PATH = os.path.join(os.path.dirname(_HERE), "include", "radnerf_b200.h")

_lib = None


class RadNerfB200Error(RuntimeError):
    pass


def declared_symbols() -> list[str]:
    """Every `rnb_*` function declared in include/radnerf_b200.h."""
    with open(HEADER_PATH) as f:
        src = f.read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(rnb_[A-Za-z0-9_]+)\s*\(", src)))


def lib() -> ctypes.CDLL:
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise RadNerfB200Error(
                f"{LIB_PATH} not built: run `python -c 'import __graft_entry__ as g; g.build()'` "
                "(make -C rad-nerf_b200/csrc). There is no CPU fallback.")
        _lib = ctypes.CDLL(LIB_PATH)
        _lib.rnb_last_error.restype = ctypes.c_char_p
        for name in declared_symbols():
            fn = getattr(_lib, name)  # raises AttributeError if the header and the .so disagree
            if name not in ("rnb_last_error",):
                fn.restype = ctypes.c_int
        _lib.rnb_engine_profile_read.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int]
    return _lib


def _arg(a):
    """tensor -> device pointer; None -> NULL; python scalars pass through typed."""
    if a is None:
        return ctypes.c_void_p(0)
    if isinstance(a, torch.Tensor):
        return ctypes.c_void_p(a.data_ptr())
    return a


def i64(v):
    return ctypes.c_int64(int(v))


def i32(v):
    return ctypes.c_int(int(v))


def f32(v):
    return ctypes.c_float(float(v))


def stream_arg(device=None):
    return ctypes.c_void_p(torch.cuda.current_stream(device).cuda_stream)


# kernels launched per C-ABI call (for bench.py's `gpu_launches` claim)
KERNELS_PER_CALL = {"rnb_raymarching_train_count": 2, "rnb_engine_gate": 3, "rnb_engine_fwd_bwd": 9, "rnb_engine_fwd_bwd_ex": 8, "rnb_engine_gate_bwd": 1, "rnb_engine_adam": 1, "rnb_engine_adam_guarded": 1, "rnb_engine_grad_check": 1,
                    "rnb_engine_layout_of": 0, "rnb_grid_update_layout_of": 0, "rnb_render_layout_of": 0,
                    "rnb_grid_update": 5, "rnb_engine_march": 3}
launch_count = 0
_prof = None


def profile_begin():
    """Start recording a CUDA-event pair around every C-ABI call (bench.py roofline leg)."""
    global _prof
    _prof = []


def profile_end():
    """-> {entry point: (calls, total_ms)} for the calls since profile_begin()."""
    global _prof
    rec, _prof = _prof, None
    torch.cuda.synchronize()
    out = {}
    for name, e0, e1 in rec or []:
        c, t = out.get(name, (0, 0.0))
        out[name] = (c + 1, t + e0.elapsed_time(e1))
    return out


def call(name: str, *args):
    """Call `rnb_<name>`; raises RadNerfB200Error with the library's message on failure."""
    global launch_count
    fn = getattr(lib(), name)
    launch_count += KERNELS_PER_CALL.get(name, 1)
    if _prof is not None:
        e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
        e0.record()
        rc = fn(*[_arg(a) for a in args])
        e1.record()
        _prof.append((name, e0, e1))
    else:
        rc = fn(*[_arg(a) for a in args])
    if rc != 0:
        msg = lib().rnb_last_error()
        raise RadNerfB200Error(f"{name} failed ({rc}): {msg.decode() if msg else ''}")


def check_input(x: torch.Tensor, name: str, dtype=None):
    """Same contract as the reference's CHECK_INPUT (models/csrc/include/utils.h:4-6):
    CUDA + contiguous, else RuntimeError; plus a dtype check the reference lacks."""
    if not x.is_cuda:
        raise RuntimeError(f"{name} must be a CUDA tensor")
    if not x.is_contiguous():
        raise RuntimeError(f"{name} must be contiguous")
    if dtype is not None and x.dtype != dtype:
        raise RuntimeError(f"{name} must be {dtype}, got {x.dtype}")
