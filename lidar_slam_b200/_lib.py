"""ctypes binding of libb200icp.so (the C ABI in include/b200icp.h).

There is deliberately NO fallback: if the library is missing or cannot be
loaded, using the engine raises.  Nothing here touches oracle/.
"""
from __future__ import annotations

import ctypes
import os

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "libb200icp.so")

c_int, c_double, c_size_t, c_void_p = ctypes.c_int, ctypes.c_double, ctypes.c_size_t, ctypes.c_void_p

# name -> (restype, argtypes); mirrors include/b200icp.h exactly
SIGNATURES = {
    "b200icp_version": (c_int, []),
    "b200icp_error_string": (ctypes.c_char_p, [c_int]),
    "b200icp_workspace_bytes": (c_int, [c_int, c_int, c_int, ctypes.POINTER(c_size_t)]),
    "b200icp_run": (c_int, [c_void_p, c_void_p, c_int, c_int, c_int, c_void_p, c_void_p, c_void_p,
                            c_void_p, c_int, c_int, c_double, c_void_p, c_void_p, c_void_p, c_void_p,
                            c_void_p, c_void_p, c_size_t, c_void_p]),
    "b200icp_host_alloc": (c_void_p, [c_size_t]),
    "b200icp_host_free": (c_int, [c_void_p]),
    "b200icp_run_static": (c_int, [c_void_p, c_void_p, c_int, c_int, c_int, c_void_p, c_void_p, c_void_p,
                                   c_void_p, c_int, c_int, c_double, c_void_p, c_void_p, c_void_p, c_void_p,
                                   c_void_p, c_void_p]),
    "b200icp_run_exchange": (c_int, [c_void_p, c_void_p, c_int, c_int, c_int, c_void_p, c_void_p, c_void_p,
                                     c_void_p, c_int, c_int, c_double, c_void_p, c_void_p, c_void_p, c_void_p,
                                     c_void_p, c_void_p, c_size_t, c_void_p, c_void_p, c_int,
                                     ctypes.c_longlong]),
    "b200icp_run_gated": (c_int, [c_void_p, c_void_p, c_int, c_int, c_int, c_void_p, c_void_p, c_void_p,
                                  c_void_p, c_int, c_int, c_double, c_void_p, c_void_p, c_void_p, c_void_p,
                                  c_void_p, c_void_p, c_size_t, c_void_p, c_void_p, c_void_p, c_int,
                                  ctypes.c_longlong]),
    "b200icp_run_ex": (c_int, [c_void_p, c_void_p, c_int, c_int, c_int, c_void_p, c_void_p, c_void_p,
                               c_void_p, c_int, c_int, c_double, c_void_p, c_void_p, c_void_p, c_void_p,
                               c_void_p, c_void_p, c_size_t, c_void_p, c_void_p, c_void_p, ctypes.c_uint]),
    "b200icp_run_ranges": (c_int, [c_void_p, c_int, c_int, c_int, c_void_p, c_void_p, c_double, c_double, c_double,
                                   c_double, c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_int, c_double,
                                   c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_size_t,
                                   c_void_p, c_void_p, c_void_p, ctypes.c_uint]),
    "b200icp_run_host_ex": (c_int, [c_void_p, c_void_p, c_int, c_int, c_int, c_void_p, c_void_p, c_void_p,
                                    c_void_p, c_int, c_int, c_double, c_void_p, c_void_p, c_void_p, c_void_p,
                                    c_void_p, c_void_p, ctypes.c_uint]),
    "b200icp_run_host_ranges": (c_int, [c_void_p, c_int, c_int, c_int, c_void_p, c_void_p, c_double, c_double,
                                        c_double, c_double, c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_int,
                                        c_double, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p,
                                        c_void_p, ctypes.c_uint]),
    "b200icp_submit_host_ranges": (c_int, [c_void_p, c_int, c_int, c_int, c_void_p, c_void_p, c_double, c_double,
                                        c_double, c_double, c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_int,
                                        c_double, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p,
                                        c_void_p, ctypes.c_uint, ctypes.POINTER(c_int)]),
    "b200icp_wait_host": (c_int, [c_int]),
    "b200icp_exchange_wait": (c_int, [c_void_p, c_int, ctypes.c_uint, c_int, c_void_p, c_void_p]),
    "b200icp_exchange_signal": (c_int, [c_void_p, c_void_p]),
    "b200icp_run_host": (c_int, [c_void_p, c_void_p, c_int, c_int, c_int, c_void_p, c_void_p, c_void_p,
                                 c_void_p, c_int, c_int, c_double, c_void_p, c_void_p, c_void_p, c_void_p,
                                 c_void_p]),
    "b200icp_run_pair_host": (c_int, [c_void_p, c_int, c_void_p, c_int, c_int, c_void_p, c_double, c_int, c_double,
                                      c_int, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p]),
    "b200icp_host_release": (c_int, []),
    "b200icp_nearest": (c_int, [c_void_p, c_void_p, c_int, c_int, c_int, c_void_p, c_int, c_void_p,
                                c_void_p, c_void_p, c_size_t, c_void_p]),
    "b200icp_pack_poses": (c_int, [c_void_p, c_void_p, c_int, c_void_p, c_void_p]),
    "b200icp_ffma_probe": (c_int, [c_void_p, c_int, c_int, c_void_p]),
    "b200icp_voxel_workspace_bytes": (c_int, [c_int, ctypes.POINTER(c_size_t)]),
    "b200icp_voxel_downsample": (c_int, [c_void_p, c_int, c_double, c_void_p, c_void_p, c_void_p,
                                         c_size_t, c_void_p]),
    "b200icp_scans_from_ranges": (c_int, [c_void_p, c_int, c_int, c_void_p, c_void_p, c_double, c_double,
                                          c_double, c_double, c_void_p, c_void_p, c_int, c_void_p]),
    "b200icp_range_filter_workspace_bytes": (c_int, [ctypes.POINTER(c_size_t)]),
    "b200icp_range_filter": (c_int, [c_void_p, c_void_p, c_int, c_int, c_double, c_void_p, c_void_p,
                                     c_size_t, c_void_p]),
    "b200icp_ogm_workspace_bytes": (c_int, [c_int, c_int, c_int, c_int, ctypes.POINTER(c_size_t)]),
    "b200icp_ogm_build": (c_int, [c_void_p, c_int, c_int, c_double, c_double, c_double, c_void_p, c_void_p,
                                  c_void_p, c_void_p, c_int, c_int, c_int, c_void_p, c_size_t, c_void_p]),
    "b200icp_ogm_binarize": (c_int, [c_void_p, c_void_p, ctypes.c_longlong, ctypes.c_float, c_void_p]),
    "b200icp_texture_paint": (c_int, [c_void_p, c_int, c_int, c_double, c_double, c_double, c_void_p, c_void_p,
                                      c_int, c_int, c_int, c_int, c_void_p, c_void_p, c_void_p, c_void_p,
                                      c_void_p]),
    "b200icp_texture_resolve": (c_int, [c_void_p, c_void_p, c_void_p, c_int, c_int, c_void_p]),
}

# launch flags (include/b200icp.h)
STATIC_SCRATCH_BYTES = 16 << 20      # B200ICP_STATIC_SCRATCH_BYTES (b200icp_run_static)
ERR_STATIC_SCRATCH = -6
FLAG_EXHAUSTIVE = 0x1
FLAG_HOST_CHUNKED = 0x2
FLAG_NO_PRUNE = 0x4
FLAG_LATENCY = 0x8
FLAG_HOST_NO_DIRECT = 0x10


def flag_slices(n: int) -> int:
    return (int(n) & 0xff) << 8


def flag_chunks(n: int) -> int:
    return (int(n) & 0xff) << 16


class Exchange(ctypes.Structure):
    """b200icp_exchange (include/b200icp.h)."""
    _fields_ = [("peer_pose_out", c_void_p), ("n_peers", c_int), ("peer_row0", ctypes.c_longlong),
                ("flag_offset_bytes", ctypes.c_longlong), ("flag_index", c_int), ("flag_value", ctypes.c_uint)]


def make_exchange(peer_pose_ptrs, peer_row0=0, flag_offset_bytes=0, flag_index=0, flag_value=0):
    """(Exchange, keep-alive) for a list of peer-mapped device pointers; None when there are no peers."""
    if not peer_pose_ptrs:
        return None, None
    arr = (c_void_p * len(peer_pose_ptrs))(*[int(q) for q in peer_pose_ptrs])
    ex = Exchange(ctypes.cast(arr, c_void_p), len(peer_pose_ptrs), int(peer_row0), int(flag_offset_bytes),
                  int(flag_index), int(flag_value))
    return ex, arr


_lib = None


def load():
    """Load the shared library (once).  Raises RuntimeError if it is not built."""
    global _lib
    if _lib is not None:
        return _lib
    global LIB_PATH
    if os.environ.get("B200ICP_LIB"):          # developer A/B: another build of the same library
        LIB_PATH = os.environ["B200ICP_LIB"]
    if not os.path.exists(LIB_PATH):
        raise RuntimeError(
            f"{LIB_PATH} is missing: build it with `python -m lidar_slam_b200.build` "
            "(nvcc, sm_100a).  There is no CPU fallback.")
    lib = ctypes.CDLL(LIB_PATH)
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(lib, name)          # AttributeError if a declared symbol is not exported
        fn.restype = res
        fn.argtypes = args
    _lib = lib
    return lib


def check(code: int, what: str):
    if code != 0:
        msg = load().b200icp_error_string(code).decode()
        if code in (-4, -5):
            msg += " -- lidar_slam_b200 needs a CUDA device (B200, sm_100a); there is no CPU fallback"
        raise RuntimeError(f"{what} failed: {msg} (code {code})")
