"""ctypes loaders for the native libraries (built in-tree by ``_build.py``)."""
from __future__ import annotations

import ctypes as C
import os
import threading
from pathlib import Path

_LIB_DIR = Path(__file__).resolve().parent / "_lib"
_lock = threading.Lock()
_host = None
_cuda = None

c_i64, c_i32, c_vp, c_cp, c_dbl = C.c_int64, C.c_int32, C.c_void_p, C.c_char_p, C.c_double


def host_lib() -> C.CDLL:
    """libdcp_host.so (store, mesh, FD passing, reducer).  Built on first use if missing."""
    global _host
    with _lock:
        if _host is not None:
            return _host
        path = _LIB_DIR / "libdcp_host.so"
        if not path.exists() or os.environ.get("DCP_REBUILD"):
            from . import _build
            _build.build_host()
        lib = C.CDLL(str(path))
        sig = {
            "dcp_store_server_start": (c_vp, [c_cp, C.c_int, C.POINTER(C.c_int)]),
            "dcp_store_server_stop": (None, [c_vp]),
            "dcp_store_client_connect": (c_vp, [c_cp, C.c_int, c_dbl]),
            "dcp_store_client_close": (None, [c_vp]),
            "dcp_store_set": (c_i64, [c_vp, c_cp, c_vp, c_i64]),
            "dcp_store_get": (c_i64, [c_vp, c_cp, c_dbl]),
            "dcp_store_copy_last": (c_i64, [c_vp, c_vp, c_i64]),
            "dcp_store_add": (c_i64, [c_vp, c_cp, c_i64]),
            "dcp_store_check": (c_i64, [c_vp, c_cp]),
            "dcp_store_delete": (c_i64, [c_vp, c_cp]),
            "dcp_store_num_keys": (c_i64, [c_vp]),
            "dcp_mesh_create": (c_vp, [c_vp, C.c_int, C.c_int, c_cp, c_cp, c_dbl]),
            "dcp_mesh_destroy": (None, [c_vp]),
            "dcp_mesh_send": (C.c_int, [c_vp, C.c_int, c_vp, c_i64, C.c_int]),
            "dcp_mesh_recv": (C.c_int, [c_vp, C.c_int, c_vp, c_i64, C.c_int]),
            "dcp_mesh_sendrecv": (C.c_int, [c_vp, C.c_int, c_vp, c_i64, C.c_int, c_vp, c_i64, C.c_int]),
            "dcp_fd_listen": (C.c_int, [c_cp]),
            "dcp_fd_send": (C.c_int, [c_cp, C.c_int, C.POINTER(C.c_int), C.c_int, c_dbl]),
            "dcp_fd_recv": (C.c_int, [C.c_int, C.POINTER(C.c_int), C.c_int, C.POINTER(C.c_int), c_dbl]),
            "dcp_fd_close": (None, [C.c_int]),
            "dcp_bucket_assign": (C.c_int, [C.POINTER(c_i64), C.POINTER(c_i32), C.c_int, C.POINTER(c_i64), C.c_int, C.POINTER(c_i32)]),
            "dcp_reducer_create": (c_vp, [C.POINTER(c_i32), C.c_int, C.c_int]),
            "dcp_reducer_destroy": (None, [c_vp]),
            "dcp_reducer_reset": (None, [c_vp]),
            "dcp_reducer_mark_ready": (C.c_int, [c_vp, C.c_int, C.POINTER(C.c_int)]),
            "dcp_reducer_unlaunched": (C.c_int, [c_vp]),
            "dcp_reducer_ready_order": (C.c_int, [c_vp, C.POINTER(c_i32), C.c_int]),
        }
        for name, (res, args) in sig.items():
            fn = getattr(lib, name)
            fn.restype, fn.argtypes = res, args
        _host = lib
        return lib


def cuda_lib_path() -> Path:
    return _LIB_DIR / "libdcp_cuda.so"


def cuda_lib() -> C.CDLL:
    """libdcp_cuda.so (sm_100a kernels).  Raises if it has not been built: on a GPU box a missing
    extension must be loud, never a silent fallback to library kernels."""
    global _cuda
    with _lock:
        if _cuda is not None:
            return _cuda
        path = cuda_lib_path()
        if not path.exists():
            from . import _build
            if _build.build_cuda() is None or not path.exists():
                raise RuntimeError(
                    f"{path} is missing and nvcc is unavailable: run `python -m distributed_compute_pytorch_b200._build`")
        lib = C.CDLL(str(path))
        from ._cuda_sigs import declare
        declare(lib)
        _cuda = lib
        return lib
