"""ctypes signatures of libdcp_cuda.so (kept in one place; mirrors include/dcp_b200.h)."""
import ctypes as C

vp, i32, i64, sz, f32, f64, cp = C.c_void_p, C.c_int, C.c_int64, C.c_size_t, C.c_float, C.c_double, C.c_char_p

SIGS = {
    "dcp_last_error": (cp, []),
    # symmetric memory
    "dcp_symm_multicast_supported": (i32, [i32]),
    "dcp_symm_create": (vp, [i32, i32, i32, sz, i32]),
    "dcp_symm_bytes": (sz, [vp]),
    "dcp_symm_local_ptr": (vp, [vp]),
    "dcp_symm_peer_ptr": (vp, [vp, i32]),
    "dcp_symm_mc_ptr": (vp, [vp]),
    "dcp_symm_export_fd": (i32, [vp]),
    "dcp_symm_import_fd": (i32, [vp, i32, i32]),
    "dcp_symm_ipc_handle": (i32, [vp, vp]),
    "dcp_symm_ipc_open": (i32, [vp, i32, vp]),
    "dcp_symm_mc_create": (i32, [vp]),
    "dcp_symm_mc_import": (i32, [vp, i32]),
    "dcp_symm_mc_add_device": (i32, [vp]),
    "dcp_symm_mc_bind": (i32, [vp]),
    "dcp_symm_finalize": (i32, [vp, f64, sz, sz, i32]),
    "dcp_symm_user_offset": (sz, [vp]),
    "dcp_symm_staging_offset": (sz, [vp]),
    "dcp_symm_staging_bytes": (sz, [vp]),
    "dcp_symm_has_nvls": (i32, [vp]),
    "dcp_symm_error_flag": (i32, [vp]),
    "dcp_symm_clear_error": (None, [vp]),
    "dcp_symm_destroy": (None, [vp]),
    # collectives
    "dcp_coll_oneshot_max_bytes": (sz, [vp]),
    "dcp_coll_bounce_max_bytes": (sz, [vp]),
    "dcp_all_reduce": (i32, [vp, vp, sz, i32, i32, i32, i32, i32, vp]),
    "dcp_broadcast": (i32, [vp, vp, sz, i32, i32, i32, vp]),
    "dcp_all_gather": (i32, [vp, vp, vp, sz, i32, i32, i32, vp]),
    "dcp_scatter": (i32, [vp, vp, vp, sz, i32, i32, i32, vp]),
    "dcp_reduce_scatter": (i32, [vp, vp, vp, sz, i32, i32, i32, i32, i32, vp]),
    "dcp_send": (i32, [vp, vp, sz, i32, vp]),
    "dcp_recv": (i32, [vp, vp, sz, i32, vp]),
    "dcp_barrier_launch": (i32, [vp, i32, vp]),
}


def declare(lib) -> None:
    for name, (res, args) in SIGS.items():
        try:
            fn = getattr(lib, name)
        except AttributeError as e:  # a stale .so: fail loudly
            raise RuntimeError(f"libdcp_cuda.so lacks symbol {name}; rebuild with `python -m distributed_compute_pytorch_b200._build --force`") from e
        fn.restype, fn.argtypes = res, args
