"""ctypes binding of libflatip_b200.so (C ABI: include/flatip_b200.h).

There is deliberately NO fallback: if the CUDA library is missing or no sm_100 device is visible,
every entry point raises.  PyTorch is optional here -- it is only used by callers that want to
hand device tensors across (``data_ptr()``), never for the arithmetic.
"""

from __future__ import annotations

import ctypes
import os
import threading

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_NAME = "libflatip_b200.so"
# FLATIP_B200_LIBRARY: developer override (A/B builds of the kernels)
LIB_PATH = os.environ.get("FLATIP_B200_LIBRARY") or os.path.join(_HERE, LIB_NAME)

FIP_F32, FIP_F16, FIP_BF16 = 0, 1, 2
DTYPE_CODES = {"f32": FIP_F32, "f16": FIP_F16, "bf16": FIP_BF16,
               "float32": FIP_F32, "float16": FIP_F16, "bfloat16": FIP_BF16}
DTYPE_NAMES = {FIP_F32: "f32", FIP_F16: "f16", FIP_BF16: "bf16"}
VARIANTS = {"auto": 0, "stream": 1, "tensor": 2}
VARIANT_NAMES = {0: "none", 1: "stream", 2: "tensor"}

FIP_OK = 0
FIP_ERR_INVALID = -1
FIP_ERR_CUDA = -2
FIP_ERR_NOMEM = -3
FIP_ERR_NODEVICE = -4
FIP_ERR_STATE = -5

# Every symbol include/flatip_b200.h declares; tests check the library exports all of them.
EXPORTED_SYMBOLS = (
    "fip_create", "fip_destroy", "fip_add_host_f32", "fip_add_device", "fip_add_synthetic",
    "fip_reset", "fip_reconstruct", "fip_ntotal", "fip_dim", "fip_dtype_of", "fip_search",
    "fip_search_device", "fip_merge_device", "fip_merge_strided_device", "fip_set_option", "fip_get_stat", "fip_last_error",
    "fip_abi_version", "fip_compiled_sm",
    "fip_exchange_create", "fip_exchange_ipc_handle", "fip_exchange_connect_ipc",
    "fip_exchange_connect_local", "fip_exchange_merge", "fip_exchange_merge_status",
    "fip_exchange_destroy",
    "fip_add_host", "fip_search_device_async", "fip_search_status", "fip_search_flags_device",
    "fip_group_create", "fip_group_create_weighted", "fip_calibrate_devices", "fip_group_destroy", "fip_group_size", "fip_group_shard", "fip_group_ntotal",
    "fip_group_add_synthetic", "fip_group_add_host", "fip_group_reset", "fip_group_search",
    "fip_group_set_option", "fip_group_get_stat",
    "fip_search_rows_static",
)
# include/ivfflat_b200.h (IVF-Flat L2 operator, same shared library)
IVF_EXPORTED_SYMBOLS = (
    "fivf_create", "fivf_destroy", "fivf_ntotal", "fivf_nlist", "fivf_dim", "fivf_is_trained",
    "fivf_train", "fivf_set_centroids", "fivf_get_centroids", "fivf_add", "fivf_add_preassigned",
    "fivf_assign", "fivf_list_sizes", "fivf_export_lists", "fivf_search", "fivf_search_device",
    "fivf_set_option", "fivf_get_stat",
)
IPC_HANDLE_BYTES = 64


class FipConfig(ctypes.Structure):
    _fields_ = [("dim", ctypes.c_int32), ("dtype", ctypes.c_int32), ("device", ctypes.c_int32),
                ("reserved", ctypes.c_int32), ("capacity_rows", ctypes.c_int64),
                ("id_base", ctypes.c_int64)]


class NativeLibraryError(RuntimeError):
    """The CUDA library is missing or cannot be loaded (no CPU fallback exists)."""


class FlatIPError(RuntimeError):
    def __init__(self, code: int, message: str):
        super().__init__(f"flatip_b200 error {code}: {message}")
        self.code = code


_lib = None
_lock = threading.Lock()


def load_library():
    """Load libflatip_b200.so; raises NativeLibraryError when it has not been built."""
    global _lib
    with _lock:
        if _lib is not None:
            return _lib
        if not os.path.exists(LIB_PATH):
            raise NativeLibraryError(
                f"{LIB_PATH} not found. Build it with `python -c 'import __graft_entry__ as g; "
                f"g.build()'` or `make -C {os.path.join(_HERE, 'csrc')}`. "
                "There is no CPU fallback for the flat-IP search path.")
        try:
            lib = ctypes.CDLL(LIB_PATH)
        except OSError as e:  # pragma: no cover - depends on the box
            raise NativeLibraryError(f"cannot load {LIB_PATH}: {e}") from e
        c = ctypes
        P = c.c_void_p
        lib.fip_create.argtypes = [c.POINTER(FipConfig), c.POINTER(P)]
        lib.fip_create.restype = c.c_int
        lib.fip_destroy.argtypes = [P]
        lib.fip_destroy.restype = None
        lib.fip_add_host_f32.argtypes = [P, P, c.c_int64]
        lib.fip_add_host_f32.restype = c.c_int
        lib.fip_add_device.argtypes = [P, P, c.c_int64, c.c_int32]
        lib.fip_add_device.restype = c.c_int
        lib.fip_add_synthetic.argtypes = [P, c.c_int64, c.c_uint64, c.c_int64]
        lib.fip_add_synthetic.restype = c.c_int
        lib.fip_reset.argtypes = [P]
        lib.fip_reset.restype = c.c_int
        lib.fip_reconstruct.argtypes = [P, c.c_int64, c.c_int64, P]
        lib.fip_reconstruct.restype = c.c_int
        lib.fip_ntotal.argtypes = [P]
        lib.fip_ntotal.restype = c.c_int64
        lib.fip_dim.argtypes = [P]
        lib.fip_dim.restype = c.c_int32
        lib.fip_dtype_of.argtypes = [P]
        lib.fip_dtype_of.restype = c.c_int32
        lib.fip_search.argtypes = [P, P, c.c_int64, c.c_int32, P, P]
        lib.fip_search.restype = c.c_int
        lib.fip_search_device.argtypes = [P, P, c.c_int64, c.c_int32, P, P, P]
        lib.fip_search_device.restype = c.c_int
        lib.fip_search_rows_static.argtypes = [c.c_int32, P, c.c_int64, c.c_int32, c.c_int32, c.c_int64, P,
                                               c.c_int64, c.c_int32, P, P, P]
        lib.fip_search_rows_static.restype = c.c_int
        lib.fip_merge_device.argtypes = [c.c_int32, P, P, c.c_int32, c.c_int64, c.c_int32,
                                         c.c_int32, P, P, P]
        lib.fip_merge_device.restype = c.c_int
        lib.fip_merge_strided_device.argtypes = [c.c_int32, P, P, c.c_int64, c.c_int64, c.c_int32,
                                                 c.c_int64, c.c_int32, c.c_int32, P, P, P]
        lib.fip_merge_strided_device.restype = c.c_int
        lib.fip_exchange_create.argtypes = [c.c_int32, c.c_int32, c.c_int32, c.c_int64, c.c_int32,
                                            c.POINTER(P)]
        lib.fip_exchange_create.restype = c.c_int
        lib.fip_exchange_ipc_handle.argtypes = [P, P]
        lib.fip_exchange_ipc_handle.restype = c.c_int
        lib.fip_exchange_connect_ipc.argtypes = [P, P]
        lib.fip_exchange_connect_ipc.restype = c.c_int
        lib.fip_exchange_connect_local.argtypes = [P, c.POINTER(P)]
        lib.fip_exchange_connect_local.restype = c.c_int
        lib.fip_exchange_merge.argtypes = [P, P, P, c.c_int64, c.c_int32, P, P, P]
        lib.fip_exchange_merge.restype = c.c_int
        lib.fip_exchange_destroy.argtypes = [P]
        lib.fip_exchange_destroy.restype = None
        lib.fip_exchange_merge_status.argtypes = [P, P, P, c.c_int64, c.c_int32, P, P, P, P, P]
        lib.fip_exchange_merge_status.restype = c.c_int
        lib.fip_add_host.argtypes = [P, P, c.c_int64, c.c_int32]
        lib.fip_add_host.restype = c.c_int
        lib.fip_search_device_async.argtypes = [P, P, c.c_int64, c.c_int32, P, P, P]
        lib.fip_search_device_async.restype = c.c_int
        lib.fip_search_status.argtypes = [P, c.POINTER(c.c_int32)]
        lib.fip_search_status.restype = c.c_int
        lib.fip_search_flags_device.argtypes = [P]
        lib.fip_search_flags_device.restype = P
        lib.fip_group_create.argtypes = [c.POINTER(FipConfig), c.POINTER(c.c_int32), c.c_int32, c.c_int64,
                                         c.POINTER(P)]
        lib.fip_group_create.restype = c.c_int
        lib.fip_group_create_weighted.argtypes = [c.POINTER(FipConfig), c.POINTER(c.c_int32), c.c_int32,
                                                  c.c_int64, c.POINTER(c.c_double), c.POINTER(P)]
        lib.fip_group_create_weighted.restype = c.c_int
        lib.fip_calibrate_devices.argtypes = [c.POINTER(c.c_int32), c.c_int32, c.c_int32, c.c_int32, c.c_int64,
                                              c.c_int32, c.c_int32, c.c_double, c.POINTER(c.c_double)]
        lib.fip_calibrate_devices.restype = c.c_int
        lib.fip_group_destroy.argtypes = [P]
        lib.fip_group_destroy.restype = None
        lib.fip_group_size.argtypes = [P]
        lib.fip_group_size.restype = c.c_int32
        lib.fip_group_shard.argtypes = [P, c.c_int32]
        lib.fip_group_shard.restype = P
        lib.fip_group_ntotal.argtypes = [P]
        lib.fip_group_ntotal.restype = c.c_int64
        lib.fip_group_add_synthetic.argtypes = [P, c.c_int64, c.c_uint64]
        lib.fip_group_add_synthetic.restype = c.c_int
        lib.fip_group_add_host.argtypes = [P, P, c.c_int64, c.c_int32]
        lib.fip_group_add_host.restype = c.c_int
        lib.fip_group_reset.argtypes = [P]
        lib.fip_group_reset.restype = c.c_int
        lib.fip_group_search.argtypes = [P, P, c.c_int64, c.c_int32, P, P]
        lib.fip_group_search.restype = c.c_int
        lib.fip_group_set_option.argtypes = [P, c.c_char_p, c.c_int64]
        lib.fip_group_set_option.restype = c.c_int
        lib.fip_group_get_stat.argtypes = [P, c.c_char_p]
        lib.fip_group_get_stat.restype = c.c_int64
        lib.fip_set_option.argtypes = [P, c.c_char_p, c.c_int64]
        lib.fip_set_option.restype = c.c_int
        lib.fip_get_stat.argtypes = [P, c.c_char_p]
        lib.fip_get_stat.restype = c.c_int64
        lib.fip_last_error.argtypes = []
        lib.fip_last_error.restype = c.c_char_p
        lib.fip_abi_version.restype = c.c_int32
        # IVF-Flat operator
        lib.fivf_create.argtypes = [c.c_int32, c.c_int32, c.c_int32, c.POINTER(P)]
        lib.fivf_create.restype = c.c_int
        lib.fivf_destroy.argtypes = [P]
        lib.fivf_destroy.restype = None
        lib.fivf_ntotal.argtypes = [P]
        lib.fivf_ntotal.restype = c.c_int64
        for name in ("fivf_nlist", "fivf_dim", "fivf_is_trained"):
            getattr(lib, name).argtypes = [P]
            getattr(lib, name).restype = c.c_int32
        lib.fivf_train.argtypes = [P, P, c.c_int64, c.c_int32, c.c_uint64]
        lib.fivf_train.restype = c.c_int
        lib.fivf_set_centroids.argtypes = [P, P]
        lib.fivf_set_centroids.restype = c.c_int
        lib.fivf_get_centroids.argtypes = [P, P]
        lib.fivf_get_centroids.restype = c.c_int
        lib.fivf_add.argtypes = [P, P, c.c_int64, P]
        lib.fivf_add.restype = c.c_int
        lib.fivf_add_preassigned.argtypes = [P, P, c.c_int64, P, P]
        lib.fivf_add_preassigned.restype = c.c_int
        lib.fivf_assign.argtypes = [P, P, c.c_int64, P]
        lib.fivf_assign.restype = c.c_int
        lib.fivf_list_sizes.argtypes = [P, P]
        lib.fivf_list_sizes.restype = c.c_int
        lib.fivf_export_lists.argtypes = [P, P, P]
        lib.fivf_export_lists.restype = c.c_int
        lib.fivf_search.argtypes = [P, P, c.c_int64, c.c_int32, c.c_int32, P, P]
        lib.fivf_search.restype = c.c_int
        lib.fivf_search_device.argtypes = [P, P, c.c_int64, c.c_int32, c.c_int32, P, P]
        lib.fivf_search_device.restype = c.c_int
        lib.fivf_set_option.argtypes = [P, c.c_char_p, c.c_int64]
        lib.fivf_set_option.restype = c.c_int
        lib.fivf_get_stat.argtypes = [P, c.c_char_p]
        lib.fivf_get_stat.restype = c.c_int64
        lib.fip_compiled_sm.restype = c.c_int32
        _lib = lib
        return lib


def check(rc: int) -> None:
    if rc != FIP_OK:
        msg = load_library().fip_last_error()
        raise FlatIPError(rc, msg.decode("utf-8", "replace") if msg else "unknown error")
