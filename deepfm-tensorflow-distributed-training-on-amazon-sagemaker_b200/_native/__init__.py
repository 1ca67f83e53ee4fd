"""ctypes bindings for the in-tree native libraries.

``cuda()`` returns libdfm_cuda.so (sm_100a kernels, C ABI in include/deepfm_b200.h);
``io()`` returns libdfm_io.so (host data layer, C ABI in include/deepfm_io.h);
``cpu()`` returns libdfm_cpu.so (host compute ops of the CPU trainers, csrc/cpu/).
All are built by scripts/build_native.py; a missing library is built on first use
when a compiler is present, and is otherwise a hard error -- there is no silent
fallback to another implementation.
"""
from __future__ import annotations

import ctypes
import os
import subprocess
import sys
import threading

_HERE = os.path.dirname(os.path.abspath(__file__))
_REPO = os.path.dirname(os.path.dirname(_HERE))
_lock = threading.Lock()
_libs: dict = {}


class NativeError(RuntimeError):
    pass


def _build(which: str) -> None:
    script = os.path.join(_REPO, "scripts", "build_native.py")
    subprocess.run([sys.executable, script, which], check=True)


def _load(name: str, which: str) -> ctypes.CDLL:
    with _lock:
        if name in _libs:
            return _libs[name]
        path = os.path.join(_HERE, name)
        if not os.path.exists(path):
            _build(which)
        try:
            lib = ctypes.CDLL(path, mode=ctypes.RTLD_GLOBAL)
        except OSError as e:  # pragma: no cover
            raise NativeError(f"cannot load {path}: {e}") from e
        _libs[name] = lib
        return lib


def cuda() -> ctypes.CDLL:
    lib = _load("libdfm_cuda.so", "cuda")
    if not getattr(lib, "_dfm_typed", False):
        lib.dfm_last_error.restype = ctypes.c_char_p
        lib._dfm_typed = True
    return lib


def io() -> ctypes.CDLL:
    lib = _load("libdfm_io.so", "io")
    if not getattr(lib, "_dfm_typed", False):
        lib.dfio_last_error.restype = ctypes.c_char_p
        lib._dfm_typed = True
    return lib


def cpu() -> ctypes.CDLL:
    return _load("libdfm_cpu.so", "cpu")


def check(rc: int, what: str = "") -> None:
    """Raises NativeError with the library's message when a C entry point failed."""
    if rc != 0:
        msg = cuda().dfm_last_error().decode("utf-8", "replace")
        raise NativeError(f"{what or 'libdfm_cuda'} failed (rc={rc}): {msg}")
