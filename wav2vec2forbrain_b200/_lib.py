"""ctypes loader for csrc/libb2t_sm100.so (the C ABI declared in include/b2t_sm100.h).

The library is the product: there is no Python/torch fallback.  Importing this module on a box
where the shared object is missing, or calling into it on a non-sm_100 device, raises.
"""
from __future__ import annotations

import ctypes
import os

import torch

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "csrc", "libb2t_sm100.so")

_lib = None
_device_ok = False

c_void_p = ctypes.c_void_p
c_int = ctypes.c_int
c_int32 = ctypes.c_int32
c_int64 = ctypes.c_int64
c_float = ctypes.c_float


class GemmArgs(ctypes.Structure):
    """Mirror of b2t_gemm_args (include/b2t_sm100.h)."""
    _fields_ = [
        ("a_hi", c_void_p), ("a_lo", c_void_p), ("b_hi", c_void_p), ("b_lo", c_void_p),
        ("a_d0", c_int64), ("a_d1", c_int64), ("a_d2", c_int64), ("a_s1", c_int64), ("a_s2", c_int64),
        ("b_d0", c_int64), ("b_d1", c_int64), ("b_d2", c_int64), ("b_s1", c_int64), ("b_s2", c_int64),
        ("a_mn_major", c_int32), ("b_mn_major", c_int32),
        ("k_spans_batches", c_int32),
        ("m_per_batch", c_int32), ("batches", c_int32), ("n", c_int32), ("k", c_int32),
        ("a_batched", c_int32), ("b_batched", c_int32),
        ("b_batch_idx", c_void_p), ("c_batch_idx", c_void_p),
        ("out_f32", c_void_p), ("out_pre", c_void_p), ("out_hi", c_void_p), ("out_lo", c_void_p),
        ("ldc", c_int64), ("c_batch_stride", c_int64),
        ("bias", c_void_p), ("residual", c_void_p), ("aux", c_void_p), ("colsum", c_void_p),
        ("act", c_int32), ("aux_mode", c_int32), ("atomic", c_int32), ("split_k", c_int32),
        ("alpha", c_float), ("block_n", c_int32),
        ("drop_p", c_float), ("drop_seed", ctypes.c_uint64), ("bias_batch_stride", c_int64),
        ("out_scale", c_float),
        ("sk_workspace", c_void_p), ("sk_workspace_bytes", c_int64), ("stream_k", c_int32),
        ("cta_group", c_int32), ("pre_mode", c_int32), ("reserved_", c_int32),
    ]


def lib() -> ctypes.CDLL:
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise RuntimeError(
                f"{LIB_PATH} is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
                "(wav2vec2forbrain_b200 has no CPU or PyTorch fallback)")
        _lib = ctypes.CDLL(LIB_PATH)
        _lib.b2t_last_error.restype = ctypes.c_char_p
        _lib.b2t_abi_version.restype = c_int
        _lib.b2t_gemm_workspace_bytes.restype = c_int64
    return _lib


def ensure_device() -> None:
    """Fail loudly unless the current CUDA device is a B200-class (sm_100) part."""
    global _device_ok
    if _device_ok:
        return
    if not torch.cuda.is_available():
        raise RuntimeError("wav2vec2forbrain_b200 needs a CUDA device (sm_100a); none is visible")
    rc = lib().b2t_check_device(c_int(torch.cuda.current_device()))
    if rc != 0:
        raise RuntimeError(lib().b2t_last_error().decode())
    _device_ok = True


def check(rc: int, what: str = "") -> None:
    if rc != 0:
        msg = lib().b2t_last_error().decode()
        raise RuntimeError(f"libb2t_sm100 {what} failed (rc={rc}): {msg}")


def ptr(t) -> c_void_p:
    """Device pointer of a tensor (None -> NULL)."""
    if t is None:
        return c_void_p(0)
    return c_void_p(t.data_ptr())


def raw_stream() -> int:
    """cudaStream_t of torch's current stream on the current device.  The private C accessors cost ~0.3 us;
    torch.cuda.current_stream() builds a Stream object through several Python layers (~5 us x 1800 launches
    per conformer step was a quarter of the host time of a step)."""
    return torch._C._cuda_getCurrentRawStream(torch._C._cuda_getDevice())


def stream() -> c_void_p:
    return c_void_p(raw_stream())
