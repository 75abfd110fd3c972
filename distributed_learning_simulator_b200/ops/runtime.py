"""ctypes binding of libdls_b200.so (C ABI in include/dls_b200.h).

``available()`` is True only when CUDA is usable AND the library loads.  On a CUDA box a
missing library is an error, not a silent fallback: ``require()`` raises with the build
hint (set DLS_ALLOW_TORCH_FALLBACK=1 to opt into the torch oracles on a GPU).
"""
import ctypes
import os
import threading

import torch

_LIB_PATH = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "lib", "libdls_b200.so")
_lock = threading.Lock()
_lib: ctypes.CDLL | None = None
_load_error: str | None = None

c_void_p, c_int, c_ll, c_float, c_double = ctypes.c_void_p, ctypes.c_int, ctypes.c_longlong, ctypes.c_float, ctypes.c_double
c_u32, c_u64 = ctypes.c_uint32, ctypes.c_uint64


class DlsPackedRow(ctypes.Structure):
    _fields_ = [("payload", c_void_p), ("aux", c_void_p), ("scale", c_void_p), ("offset", c_void_p),
                ("bits", c_void_p), ("byte_offsets", c_void_p), ("present", c_void_p), ("weight", c_float),
                ("kind", ctypes.c_int32), ("is_delta", ctypes.c_int32), ("pad", ctypes.c_int32)]


_SIGNATURES = {
    "dls_fedavg_reduce_f32": [c_void_p, c_void_p, c_int, c_void_p, c_float, c_double, c_void_p, c_int, c_ll, c_ll,
                              c_void_p, c_int, c_int, c_u32, c_void_p, c_void_p],
    "dls_fedavg_reduce_f32_mc": [c_void_p, c_void_p, c_int, c_void_p, c_float, c_double, c_void_p, c_ll, c_ll,
                                 c_void_p, c_int, c_int, c_u32, c_void_p, c_void_p],
    "dls_multimem_bcast_f32": [c_void_p, c_void_p, c_ll, c_void_p],
    "dls_fedavg_reduce_packed": [c_void_p, c_int, c_void_p, c_void_p, c_void_p, c_int, c_int, c_void_p, c_void_p,
                                 c_void_p],
    "dls_sgd_step": [c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_void_p, c_int, c_void_p, c_float, c_float,
                     c_float, c_int, c_void_p, c_void_p, c_void_p],
    "dls_sgd_tick": [c_void_p, c_void_p],
    "dls_gather_batch": [c_void_p, c_void_p, c_void_p, c_ll, c_void_p, c_int, c_int, c_int, c_ll, c_int,
                         c_void_p, c_void_p, c_void_p],
    "dls_metric_accumulate": [c_void_p, c_void_p, c_void_p, c_int, c_int, c_void_p, c_void_p],
    "dls_sq_quant": [c_void_p, c_ll, c_void_p, c_int, c_int, c_u64, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p],
    "dls_sq_dequant": [c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_ll, c_void_p, c_void_p],
    "dls_nnadq_quant": [c_void_p, c_void_p, c_void_p, c_int, c_float, c_void_p, c_void_p, c_void_p, c_void_p,
                        c_void_p, c_void_p, c_ll, c_void_p],
    "dls_nnadq_dequant": [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_void_p,
                          c_void_p],
    "dls_ipc_alloc": [c_ll, ctypes.POINTER(c_void_p)],
    "dls_ipc_free": [c_void_p],
    "dls_ipc_get_handle": [c_void_p, ctypes.c_char_p],
    "dls_ipc_open_handle": [ctypes.c_char_p, ctypes.POINTER(c_void_p)],
    "dls_ipc_close_handle": [c_void_p],
    "dls_enable_peer_access": [c_int, c_int],
    "dls_copy_device": [c_void_p, c_void_p, c_ll, c_void_p],
    "dls_gemm_bf16_tn": [c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_int, c_int, c_int, c_int,
                         c_void_p],
    "dls_conv_supported": [c_void_p],
    "dls_conv_fprop": [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p],
    "dls_conv_dgrad": [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p],
    "dls_conv_wgrad": [c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_void_p],
    "dls_conv_direct_fprop": [c_void_p, c_void_p, c_void_p, c_int, c_void_p, c_void_p, c_void_p],
    "dls_conv_direct_dgrad": [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p],
    "dls_conv_direct_wgrad": [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p],
    "dls_conv_stem_im2col": [c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_void_p, c_void_p],
    "dls_tensor_map_probe": [c_int, c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_void_p],
    "dls_bn_scratch_floats": [c_ll, c_int, c_int],
    "dls_bn_fwd": [c_void_p, c_int, c_ll, c_int, c_void_p, c_void_p, c_void_p, c_float, c_int, c_void_p, c_void_p,
                   c_void_p, c_void_p, c_void_p],
    "dls_bn_fwd_stats": [c_void_p, c_int, c_ll, c_int, c_void_p, c_void_p, c_void_p, c_float, c_int, c_void_p, c_void_p,
                         c_void_p, c_void_p, c_void_p],
    "dls_bn_bwd": [c_void_p, c_void_p, c_void_p, c_int, c_ll, c_int, c_void_p, c_void_p, c_void_p, c_int, c_void_p,
                   c_void_p, c_void_p, c_void_p, c_void_p, c_void_p],
    "dls_bn_bwd_acc": [c_void_p, c_void_p, c_void_p, c_int, c_ll, c_int, c_void_p, c_void_p, c_void_p, c_int, c_void_p,
                       c_void_p, c_void_p, c_void_p, c_void_p, c_void_p],
    "dls_sign_pack": [c_void_p, c_ll, c_void_p, c_void_p],
    "dls_sign_majority": [c_void_p, c_int, c_ll, c_void_p, c_void_p],
    "dls_sign_vote_sgd": [c_void_p, c_int, c_ll, c_void_p, c_void_p, c_float, c_float, c_float, c_int, c_int, c_void_p,
                          c_void_p, c_void_p, c_int, c_void_p],
    "dls_dropout_mask": [c_void_p, c_ll, c_float, c_u64, c_void_p, c_void_p, c_void_p],
    "dls_block_sqdiff": [c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_int, c_void_p, c_void_p],
    "dls_subset_gemm_bf16": [c_void_p, c_int, c_void_p, c_int, c_ll, c_ll, c_void_p, c_int, c_int, c_int, c_void_p,
                             c_int, c_int, c_int, c_int, c_int, c_int, c_void_p],
    "dls_im2col_grouped": [c_void_p, c_int, c_int, c_int, c_int, c_int, c_int, c_int, c_int, c_int, c_void_p, c_void_p],
    "dls_im2col_grouped_padded": [c_void_p, c_int, c_int, c_int, c_int, c_int, c_int, c_int, c_int, c_int, c_int, c_void_p,
                                  c_void_p],
    "dls_index_compact": [c_void_p, c_void_p, c_ll, c_void_p, c_void_p, c_float, c_u64, c_void_p, c_void_p, c_void_p,
                          c_void_p, c_void_p, c_void_p],
    "dls_row_gather": [c_void_p, c_void_p, c_ll, c_ll, c_void_p, c_void_p],
    "dls_halo_merge": [c_void_p, c_void_p, c_ll, c_ll, c_void_p, c_void_p, c_void_p, c_ll, c_void_p, c_void_p,
                       c_void_p, c_void_p, c_void_p],
    "dls_gcn_dinv": [c_void_p, c_ll, c_ll, c_void_p, c_void_p, c_void_p],
    "dls_gcn_aggregate": [c_void_p, c_void_p, c_void_p, c_ll, c_ll, c_int, c_void_p, c_void_p, c_void_p, c_void_p],
}


def lib_path() -> str:
    return _LIB_PATH


def _load() -> ctypes.CDLL | None:
    global _lib, _load_error
    if _lib is not None or _load_error is not None:
        return _lib
    with _lock:
        if _lib is not None or _load_error is not None:
            return _lib
        if not os.path.isfile(_LIB_PATH):
            _load_error = f"{_LIB_PATH} not built (run python -m distributed_learning_simulator_b200.build)"
            return None
        try:
            lib = ctypes.CDLL(_LIB_PATH)
        except OSError as e:
            _load_error = str(e)
            return None
        lib.dls_last_error.restype = ctypes.c_char_p
        lib.dls_launch_count.restype = c_ll
        for name, argtypes in _SIGNATURES.items():
            fn = getattr(lib, name, None)
            if fn is not None:
                fn.argtypes = argtypes
                fn.restype = c_int
        _lib = lib
    return _lib


def available() -> bool:
    if not torch.cuda.is_available():
        return False
    if os.environ.get("DLS_ALLOW_TORCH_FALLBACK", "0") == "1" and _load() is None:
        return False
    return require() is not None


def require() -> ctypes.CDLL:
    lib = _load()
    if lib is None:
        raise RuntimeError(f"the CUDA kernel library is mandatory on a GPU box: {_load_error}")
    return lib


def call(name: str, *args) -> None:
    lib = require()
    rc = getattr(lib, name)(*args)
    if rc != 0:
        raise RuntimeError(f"{name} failed ({rc}): {lib.dls_last_error().decode()}")


def launch_count() -> int:
    lib = _load()
    return int(lib.dls_launch_count()) if lib is not None else 0


def reset_launch_count() -> None:
    lib = _load()
    if lib is not None:
        lib.dls_reset_launch_count()


def ptr(t: torch.Tensor | None) -> int | None:
    if t is None:
        return None
    return t.data_ptr()


def stream_ptr(device: torch.device | None = None) -> int:
    return torch.cuda.current_stream(device).cuda_stream


def pointer_table(tensors: list, device: torch.device) -> torch.Tensor:
    """Device int64 tensor holding the data pointers of ``tensors`` (None -> 0)."""
    host = torch.tensor([0 if t is None else (t if isinstance(t, int) else t.data_ptr()) for t in tensors],
                        dtype=torch.int64).pin_memory()
    return host.to(device, non_blocking=True)
