"""ctypes binding of ``libddpb200.so`` (C ABI in ``include/ddp_b200.h``).

The library is built in-tree by ``csrc/build.py`` (nvcc, sm_100a) into ``_native/``; it is loaded
lazily and *loudly*: on a machine with a GPU a missing or unloadable library is an error, never a
silent fallback to PyTorch ops.
"""
from __future__ import annotations

import ctypes
import functools
import os
import pathlib
from typing import Optional

import torch

F32, BF16 = 0, 1
P2P_ONESHOT, P2P_TWOSHOT, P2P_ROOT, P2P_BCAST, P2P_ALLGATHER = range(5)
MAX_RANKS, P2P_MAX_BLOCKS, P2P_MAX_SLOTS = 8, 320, 64

_PKG = pathlib.Path(__file__).resolve().parent.parent
LIB_PATH = _PKG / "_native" / "libddpb200.so"
REPO_ROOT = _PKG.parent


class NativeError(RuntimeError):
    pass


def dtype_code(dtype: torch.dtype) -> int:
    if dtype == torch.float32:
        return F32
    if dtype == torch.bfloat16:
        return BF16
    raise NativeError(f"unsupported dtype {dtype}")


def build(verbose: bool = False) -> pathlib.Path:
    import importlib.util
    spec = importlib.util.spec_from_file_location("_ddpb_build", REPO_ROOT / "csrc" / "build.py")
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod.build(verbose=verbose)


@functools.lru_cache(maxsize=1)
def lib() -> ctypes.CDLL:
    if not LIB_PATH.exists():
        if os.environ.get("DDPB_NO_AUTOBUILD"):
            raise NativeError(f"{LIB_PATH} is missing; run `python csrc/build.py`")
        build()
    try:
        handle = ctypes.CDLL(str(LIB_PATH))
    except OSError as e:                                   # pragma: no cover
        raise NativeError(f"cannot load {LIB_PATH}: {e}") from e
    handle.ddpb_last_error.restype = ctypes.c_char_p
    return handle


def available() -> bool:
    """True when the native engine can run here (library loadable AND a CUDA device present)."""
    if not torch.cuda.is_available():
        return False
    lib()
    return True


def check(rc: int, what: str = "") -> None:
    if rc != 0:
        msg = lib().ddpb_last_error()
        raise NativeError(f"{what or 'native call'} failed (code {rc}): {msg.decode() if msg else '?'}")


def ptr(t: Optional[torch.Tensor]) -> ctypes.c_void_p:
    if t is None:
        return ctypes.c_void_p(0)
    if not t.is_cuda:
        raise NativeError("native ops need CUDA tensors")
    if not t.is_contiguous():
        raise NativeError("native ops need contiguous tensors")
    return ctypes.c_void_p(t.data_ptr())


def stream_handle(stream: Optional[torch.cuda.Stream] = None) -> ctypes.c_void_p:
    s = stream if stream is not None else torch.cuda.current_stream()
    return ctypes.c_void_p(s.cuda_stream)


def cf(x: float) -> ctypes.c_float:
    return ctypes.c_float(x)


def cll(x: int) -> ctypes.c_longlong:
    return ctypes.c_longlong(int(x))


class DeviceBuffer:
    """A raw device allocation exposed to torch through ``__cuda_array_interface__``."""

    def __init__(self, address: int, nbytes: int) -> None:
        self.address, self.nbytes = int(address), int(nbytes)
        self.__cuda_array_interface__ = {
            "shape": (self.nbytes,), "typestr": "|u1", "data": (self.address, False), "version": 3,
            "strides": None,
        }

    def as_tensor(self, device) -> torch.Tensor:
        return torch.as_tensor(self, device=device)
