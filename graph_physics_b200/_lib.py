"""ctypes binding of libmgn_b200.so (C ABI declared in include/mgn_b200.h).

There is NO CPU or PyTorch fallback: if the library is missing, cannot be loaded,
or the current device is not an sm_100a GPU, every op raises.
"""
from __future__ import annotations

import ctypes
import os

import torch

_HERE = os.path.dirname(os.path.abspath(__file__))
# MGN_LIB_PATH: developer override (e.g. a -DMGN_TIMELINE build kept next to the production library)
LIB_PATH = os.environ.get("MGN_LIB_PATH") or os.path.join(_HERE, "libmgn_b200.so")

_lib = None
_device_ok = False


class MgnError(RuntimeError):
    pass


def lib() -> ctypes.CDLL:
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise MgnError(
                f"{LIB_PATH} not found: build it with `python -m graph_physics_b200.build` "
                "(graph_physics_b200 has no CPU / eager fallback)")
        _lib = ctypes.CDLL(LIB_PATH)
        _lib.mgn_last_error.restype = ctypes.c_char_p
        for name in ("mgn_sort_unique_workspace_bytes", "mgn_csr_workspace_bytes",
                     "mgn_khop_workspace_bytes", "mgn_filter_workspace_bytes"):
            getattr(_lib, name).restype = ctypes.c_int64
    return _lib


def require_device(t: torch.Tensor | None = None) -> None:
    """Raises unless the tensor (or the current device) is a CUDA sm_100 device."""
    global _device_ok
    if t is not None and not t.is_cuda:
        raise MgnError("graph_physics_b200 ops need CUDA tensors (no CPU fallback); got " + str(t.device))
    if not _device_ok:
        if not torch.cuda.is_available():
            raise MgnError("graph_physics_b200 needs a CUDA device (no CPU fallback)")
        check(lib().mgn_check_device())
        _device_ok = True


def hang_report() -> str:
    """One line describing the mbarrier wait that timed out inside a kernel (csrc/tc05.cuh: every wait is bounded and
    traps after writing a record to host-mapped memory), or "" if none was recorded.  Readable after the CUDA context
    has faulted."""
    if _lib is None:
        return ""
    buf = ctypes.create_string_buffer(512)
    if lib().mgn_hang_report(buf, ctypes.c_int(512)):
        return buf.value.decode("utf-8", "replace")
    return ""


def check(rc: int) -> None:
    if rc == 0:
        return
    msg = lib().mgn_last_error().decode("utf-8", "replace")
    hr = hang_report() if rc > 0 else ""
    if hr:
        msg = f"{msg} [{hr}]"
    if rc < 0:
        raise ValueError(f"libmgn_b200: {msg} (code {rc})")
    raise MgnError(f"libmgn_b200: {msg} (cudaError {rc})")


def p(t: torch.Tensor | None) -> ctypes.c_void_p:
    """Raw device pointer of a contiguous tensor (or NULL)."""
    if t is None:
        return ctypes.c_void_p(0)
    assert t.is_contiguous(), "libmgn_b200 takes contiguous buffers"
    return ctypes.c_void_p(t.data_ptr())


def i64(v: int) -> ctypes.c_int64:
    return ctypes.c_int64(int(v))


def i32(v: int) -> ctypes.c_int:
    return ctypes.c_int(int(v))


def stream() -> ctypes.c_void_p:
    return ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
