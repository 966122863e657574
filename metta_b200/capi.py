"""ctypes binding of libmettagrid_b200.so (the C-ABI declared in include/mettagrid_b200.h).

There is no fallback: if the library is missing the import error says how to build it, and any
entry point that needs a GPU returns an error code that is raised as ``MgError``.
"""

from __future__ import annotations

import ctypes
import os
from ctypes import POINTER, c_char_p, c_int, c_int32, c_int64, c_void_p

_HERE = os.path.dirname(os.path.abspath(__file__))
# MG_LIB_PATH selects another nvcc build of the same library (tuning experiments); there is no non-CUDA build.
LIB_PATH = os.environ.get("MG_LIB_PATH") or os.path.join(_HERE, "csrc", "libmettagrid_b200.so")

# every symbol include/mettagrid_b200.h declares
SYMBOLS = [
    "mg_last_error", "mg_abi_version", "mg_umap_order_u8", "mg_umap_order_str", "mg_create", "mg_destroy",
    "mg_load_maps", "mg_set_buffers", "mg_reset", "mg_step", "mg_step_host", "mg_step_host_begin", "mg_step_host_end", "mg_profile_step", "mg_compute_observations",
    "mg_set_options", "mg_set_early_reset", "mg_sizeof", "mg_read", "mg_write", "mg_set_inventory",
    "mg_kernel_launches", "mg_sync", "mg_peek_observations_noalloc", "mg_set_host_export", "mg_transfer_bytes", "mg_set_obs_aux", "mg_token_record_layout",
]  # fmt: skip


class MgError(RuntimeError):
    """A C-ABI call returned a non-zero status (the counterpart of the reference's
    std::runtime_error -> Python exception path, mettagrid_c.cpp:680-719)."""


def _declare(l: ctypes.CDLL) -> ctypes.CDLL:
    l.mg_last_error.restype = c_char_p
    l.mg_last_error.argtypes = []
    l.mg_abi_version.restype = c_int
    l.mg_umap_order_u8.argtypes = [c_void_p, c_int, c_int, c_void_p]
    l.mg_umap_order_u8.restype = c_int
    l.mg_umap_order_str.argtypes = [c_char_p, c_int, c_int, c_void_p]
    l.mg_umap_order_str.restype = c_int
    l.mg_create.argtypes = [c_void_p, c_int64, c_int32, c_int32, POINTER(c_void_p)]
    l.mg_create.restype = c_int
    l.mg_destroy.argtypes = [c_void_p]
    l.mg_destroy.restype = None
    l.mg_load_maps.argtypes = [c_void_p, c_int32] + [c_void_p] * 6
    l.mg_load_maps.restype = c_int
    l.mg_set_buffers.argtypes = [c_void_p] * 6
    l.mg_set_buffers.restype = c_int
    l.mg_reset.argtypes = [c_void_p] * 4
    l.mg_reset.restype = c_int
    l.mg_step.argtypes = [c_void_p, c_void_p]
    l.mg_step.restype = c_int
    l.mg_step_host.argtypes = [c_void_p] * 6
    l.mg_step_host.restype = c_int
    l.mg_step_host_begin.argtypes = [c_void_p] * 6
    l.mg_step_host_begin.restype = c_int
    l.mg_step_host_end.argtypes = [c_void_p]
    l.mg_step_host_end.restype = c_int
    l.mg_profile_step.argtypes = [c_void_p, c_void_p, c_void_p]
    l.mg_profile_step.restype = c_int
    l.mg_compute_observations.argtypes = [c_void_p, c_void_p]
    l.mg_compute_observations.restype = c_int
    l.mg_set_options.argtypes = [c_void_p, c_int32, c_int32]
    l.mg_set_options.restype = c_int
    l.mg_set_early_reset.argtypes = [c_void_p, c_void_p]
    l.mg_set_early_reset.restype = c_int
    l.mg_sizeof.argtypes = [c_void_p, c_int32]
    l.mg_sizeof.restype = c_int64
    l.mg_read.argtypes = [c_void_p, c_int32, c_void_p, c_int64]
    l.mg_read.restype = c_int
    l.mg_write.argtypes = [c_void_p, c_int32, c_void_p, c_int64]
    l.mg_write.restype = c_int
    l.mg_set_inventory.argtypes = [c_void_p, c_int32, c_int32, c_void_p, c_int32]
    l.mg_set_inventory.restype = c_int
    l.mg_peek_observations_noalloc.argtypes = [c_void_p, c_void_p, c_void_p]
    l.mg_peek_observations_noalloc.restype = c_int
    l.mg_set_host_export.argtypes = [c_void_p, c_int32]
    l.mg_set_host_export.restype = c_int
    l.mg_set_obs_aux.argtypes = [c_void_p] * 4
    l.mg_set_obs_aux.restype = c_int
    l.mg_token_record_layout.argtypes = [c_void_p, c_void_p]
    l.mg_token_record_layout.restype = c_int
    l.mg_transfer_bytes.argtypes = [c_void_p, c_void_p, c_void_p]
    l.mg_transfer_bytes.restype = c_int
    l.mg_kernel_launches.argtypes = [c_void_p]
    l.mg_kernel_launches.restype = c_int64
    l.mg_sync.argtypes = [c_void_p]
    l.mg_sync.restype = c_int
    return l


_lib = None


def lib() -> ctypes.CDLL:
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise ImportError(
                f"{LIB_PATH} is missing. Build it with `python -c 'import __graft_entry__ as g; g.build()'` "
                "(needs nvcc). There is no CPU implementation of the simulator to fall back to."
            )
        _lib = _declare(ctypes.CDLL(LIB_PATH))
    return _lib


def check(status: int) -> None:
    if status != 0:
        msg = lib().mg_last_error()
        raise MgError(msg.decode() if msg else f"mettagrid_b200 error {status}")
