"""ctypes binding of libnvllm_b200.so — exactly the symbols of include/nvllm.h.

This is the Python twin of the cgo shim (go/cuda_model_runner.go): plain pointers
and sizes cross the boundary, nothing else.  There is no fallback: if the shared
library is missing or no sm_100 device is usable, calls fail loudly.
"""
from __future__ import annotations

import ctypes
import os

import numpy as np

from . import config as C

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libnvllm_b200.so")

_i32p = ctypes.POINTER(ctypes.c_int32)
_f32p = ctypes.POINTER(ctypes.c_float)
_u64p = ctypes.POINTER(ctypes.c_uint64)


class RunStats(ctypes.Structure):
    """``struct nvllm_run_stats``."""

    _fields_ = [
        ("device_ms", ctypes.c_float),
        ("kernel_launches", ctypes.c_int32),
        ("new_tokens", ctypes.c_int32),
        ("used_graph", ctypes.c_int32),
        ("h2d_bytes", ctypes.c_uint64),
        ("d2h_bytes", ctypes.c_uint64),
    ]


class ProfileEntry(ctypes.Structure):
    """``struct nvllm_profile_entry``."""

    _fields_ = [("ms", ctypes.c_double), ("bytes", ctypes.c_double), ("flops", ctypes.c_double),
                ("launches", ctypes.c_int64), ("impl_bytes", ctypes.c_double)]


KERNEL_CLASSES = ["embed", "norm", "gemm_tc5", "gemv", "gemm_simt", "rope_kv", "attn_prefill",
                  "attn_decode", "act", "moe_route", "moe_gemm", "moe_misc", "sample", "misc", "comm"]


class NvllmError(RuntimeError):
    def __init__(self, code: int, msg: str):
        super().__init__(f"nvllm error {code}: {msg}")
        self.code = code


ERR_INVALID, ERR_CUDA, ERR_STATE, ERR_RANGE, ERR_NCCL, ERR_NOMEM = 1, 2, 3, 4, 5, 6

# every symbol include/nvllm.h declares: (name, restype, argtypes)
SIGNATURES = {
    "nvllm_abi_version": (ctypes.c_int, []),
    "nvllm_tp_unique_id": (ctypes.c_int, [ctypes.POINTER(ctypes.c_uint8)]),
    "nvllm_create": (ctypes.c_int, [ctypes.POINTER(C.CConfig), ctypes.POINTER(ctypes.c_uint8),
                                    ctypes.POINTER(ctypes.c_void_p)]),
    "nvllm_set_weight": (ctypes.c_int, [ctypes.c_void_p, ctypes.c_int, ctypes.c_int, ctypes.c_int,
                                        ctypes.c_void_p, ctypes.c_uint64]),
    "nvllm_finalize": (ctypes.c_int, [ctypes.c_void_p]),
    "nvllm_destroy": (ctypes.c_int, [ctypes.c_void_p]),
    "nvllm_last_error": (ctypes.c_char_p, [ctypes.c_void_p]),
    "nvllm_run": (ctypes.c_int, [ctypes.c_void_p, ctypes.c_int, ctypes.c_int, _i32p, _i32p, _i32p,
                                 _i32p, _i32p, _f32p, _f32p, _i32p, _f32p,
                                 ctypes.POINTER(RunStats)]),
    "nvllm_run_resident": (ctypes.c_int, [ctypes.c_void_p, ctypes.POINTER(RunStats)]),
    "nvllm_sync": (ctypes.c_int, [ctypes.c_void_p, _i32p, ctypes.c_int, _f32p]),
    "nvllm_rerun": (ctypes.c_int, [ctypes.c_void_p, ctypes.POINTER(RunStats)]),
    "nvllm_set_profile": (ctypes.c_int, [ctypes.c_void_p, ctypes.c_int]),
    "nvllm_get_profile": (ctypes.c_int, [ctypes.c_void_p, ctypes.POINTER(ProfileEntry)]),
    "nvllm_trace_begin": (ctypes.c_int, [ctypes.c_void_p, ctypes.c_int]),
    "nvllm_trace_end": (ctypes.c_int, [ctypes.c_void_p, _u64p, ctypes.c_int,
                                       ctypes.POINTER(ctypes.c_int)]),
    "nvllm_timer_record": (ctypes.c_int, [ctypes.c_void_p, ctypes.c_int]),
    "nvllm_timer_elapsed": (ctypes.c_int, [ctypes.c_void_p, _f32p]),
    "nvllm_set_sampling": (ctypes.c_int, [ctypes.c_void_p, ctypes.c_float, ctypes.c_int,
                                          ctypes.c_float]),
    "nvllm_set_seed": (ctypes.c_int, [ctypes.c_void_p, ctypes.c_uint64]),
    "nvllm_set_layer_types": (ctypes.c_int, [ctypes.c_void_p, ctypes.POINTER(ctypes.c_int32), ctypes.c_int]),
    "nvllm_load_safetensors": (ctypes.c_int, [ctypes.c_void_p, ctypes.c_char_p, ctypes.c_int, _u64p,
                                              ctypes.POINTER(ctypes.c_int32)]),
    "nvllm_safetensors_find": (ctypes.c_int, [ctypes.c_char_p, ctypes.c_char_p,
                                              ctypes.POINTER(ctypes.c_int32),
                                              ctypes.POINTER(ctypes.c_int32),
                                              ctypes.POINTER(ctypes.c_int64), _u64p, _u64p]),
    "nvllm_set_graphs": (ctypes.c_int, [ctypes.c_void_p, ctypes.c_int]),
    "nvllm_set_speculation": (ctypes.c_int, [ctypes.c_void_p, ctypes.c_int]),
    "nvllm_get_config": (ctypes.c_int, [ctypes.c_void_p, ctypes.POINTER(C.CConfig)]),
    "nvllm_memory_info": (ctypes.c_int, [ctypes.c_void_p, _u64p, _u64p, _u64p]),
    "nvllm_op_matmul": (ctypes.c_int, [ctypes.c_int, ctypes.c_int, ctypes.c_int, ctypes.c_int,
                                       ctypes.c_int, ctypes.c_int, _f32p, _f32p, _f32p, _f32p,
                                       ctypes.c_int, ctypes.c_int, _f32p]),
    "nvllm_op_norm": (ctypes.c_int, [ctypes.c_int, ctypes.c_int, ctypes.c_int, ctypes.c_int, _f32p,
                                     _f32p, _f32p, ctypes.c_float, _f32p, ctypes.c_int, _f32p]),
    "nvllm_op_argmax": (ctypes.c_int, [ctypes.c_int, ctypes.c_int, ctypes.c_int, _f32p, _i32p,
                                       ctypes.c_int, _f32p]),
    "nvllm_op_gemv_noalloc": (ctypes.c_int, [ctypes.c_int, ctypes.c_int, ctypes.c_int, ctypes.c_int,
                                             ctypes.c_int, ctypes.c_void_p, ctypes.c_void_p,
                                             ctypes.c_void_p, ctypes.c_void_p]),
    "nvllm_device_info": (ctypes.c_int, [ctypes.c_int, ctypes.POINTER(ctypes.c_int),
                                         ctypes.POINTER(ctypes.c_int), ctypes.POINTER(ctypes.c_int),
                                         _u64p, ctypes.c_char_p]),
}

_lib = None


def load() -> ctypes.CDLL:
    """dlopen the library and type every entry point.  No build-on-import, no fallback."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise FileNotFoundError(
                f"{LIB_PATH} is missing: run `python -c 'import __graft_entry__ as g; g.build()'` "
                "(there is no CPU fallback)")
        L = ctypes.CDLL(LIB_PATH)
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(L, name)
            fn.restype = res
            fn.argtypes = args
        if L.nvllm_abi_version() != C.ABI_VERSION:
            raise RuntimeError("libnvllm_b200.so ABI version mismatch")
        _lib = L
    return _lib


def last_error(handle=None) -> str:
    s = load().nvllm_last_error(handle)
    return s.decode() if s else ""


def check(rc: int, handle=None):
    if rc != 0:
        raise NvllmError(rc, last_error(handle))


def _fp(a):
    return a.ctypes.data_as(_f32p) if a is not None else None


def _ip(a):
    return a.ctypes.data_as(_i32p) if a is not None else None


def tp_unique_id() -> bytes:
    """128-byte NCCL unique id for a tensor-parallel group (make it on one rank, hand the same
    bytes to every rank's nvllm_create)."""
    buf = (ctypes.c_uint8 * 128)()
    check(load().nvllm_tp_unique_id(buf))
    return bytes(buf)


def device_info(device: int = 0) -> dict:
    sm, maj, mnr = ctypes.c_int(), ctypes.c_int(), ctypes.c_int()
    mem = ctypes.c_uint64()
    name = ctypes.create_string_buffer(256)
    check(load().nvllm_device_info(device, ctypes.byref(sm), ctypes.byref(maj), ctypes.byref(mnr),
                                   ctypes.byref(mem), name))
    return {"sm_count": sm.value, "cc": (maj.value, mnr.value), "total_mem": mem.value,
            "name": name.value.decode()}


def op_matmul(A, W, bias=None, prec=C.PREC_FP32, kernel=0, iters=0, flush_l2=False, device=0):
    """C = A @ W (+bias) with W [in, out] like tensor.MatMul. Returns (C, avg_ms)."""
    A = np.ascontiguousarray(A, dtype=np.float32)
    W = np.ascontiguousarray(W, dtype=np.float32)
    M, K = A.shape
    N = W.shape[1]
    out = np.empty((M, N), dtype=np.float32)
    b = np.ascontiguousarray(bias, dtype=np.float32) if bias is not None else None
    ms = ctypes.c_float()
    check(load().nvllm_op_matmul(device, prec, kernel, M, N, K, _fp(A), _fp(W), _fp(b), _fp(out),
                                 iters, int(flush_l2), ctypes.byref(ms)))
    return out, ms.value


def op_norm(x, w, b=None, eps=1e-5, prec=C.PREC_FP32, iters=0, device=0):
    x = np.ascontiguousarray(x, dtype=np.float32)
    w = np.ascontiguousarray(w, dtype=np.float32)
    bb = np.ascontiguousarray(b, dtype=np.float32) if b is not None else None
    y = np.empty_like(x)
    ms = ctypes.c_float()
    check(load().nvllm_op_norm(device, prec, x.shape[0], x.shape[1], _fp(x), _fp(w), _fp(bb), eps,
                               _fp(y), iters, ctypes.byref(ms)))
    return y, ms.value


def op_argmax(logits, iters=0, device=0):
    l = np.ascontiguousarray(logits, dtype=np.float32)
    out = np.empty((l.shape[0],), dtype=np.int32)
    ms = ctypes.c_float()
    check(load().nvllm_op_argmax(device, l.shape[0], l.shape[1], _fp(l), _ip(out), iters,
                                 ctypes.byref(ms)))
    return out, ms.value
