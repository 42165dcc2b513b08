"""ctypes binding to ``b200dp/lib/libb200dp.so`` — the C-ABI library built from
``csrc/*.cu`` (declarations: ``include/b200dp.h``)."""
from __future__ import annotations

import ctypes
import os
from typing import Optional

_LIB_PATH = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "lib", "libb200dp.so")
_lib: Optional[ctypes.CDLL] = None
_err: Optional[str] = None


def lib_path() -> str:
    return _LIB_PATH


def load() -> ctypes.CDLL:
    """Load the library or raise — GPU code paths never fall back silently."""
    global _lib, _err
    if _lib is not None:
        return _lib
    if not os.path.exists(_LIB_PATH):
        _err = f"{_LIB_PATH} not built; run `python -c 'import __graft_entry__ as g; g.build()'` or `make -C csrc`"
        raise RuntimeError(_err)
    _lib = ctypes.CDLL(_LIB_PATH, mode=ctypes.RTLD_GLOBAL)
    _declare(_lib)
    return _lib


def available() -> bool:
    try:
        load()
        return True
    except (RuntimeError, OSError):
        return False


def _declare(lib: ctypes.CDLL) -> None:
    from . import _sigs
    _sigs.declare(lib)
