"""ctypes binding of oracle/quant_oracle.c (plain-C restatement).  TEST INFRASTRUCTURE ONLY."""
from __future__ import annotations

import ctypes
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "_build", "libsglk_oracle.so")
DT = {"bf16": 0, "f16": 1, "f32": 2}


def build(force: bool = False) -> str:
    src = os.path.join(_HERE, "quant_oracle.c")
    if force or not os.path.exists(_SO) or os.path.getmtime(_SO) < os.path.getmtime(src):
        subprocess.check_call(["make", "-C", _HERE, "-s"])
    return _SO


_lib = None


def lib() -> ctypes.CDLL:
    global _lib
    if _lib is None:
        L = ctypes.CDLL(build())
        L.oracle_f32_to_e4m3fn.restype = ctypes.c_uint8
        L.oracle_f32_to_e4m3fn.argtypes = [ctypes.c_float]
        L.oracle_e4m3fn_to_f32.restype = ctypes.c_float
        L.oracle_e4m3fn_to_f32.argtypes = [ctypes.c_uint8]
        L.oracle_per_token_group_quant.restype = None
        L.oracle_per_token_group_quant.argtypes = [
            ctypes.c_void_p, ctypes.c_int, ctypes.c_int64, ctypes.c_int64, ctypes.c_int64,
            ctypes.c_float, ctypes.c_int, ctypes.c_int, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p]
        L.oracle_blockwise_scaled_mm.restype = None
        L.oracle_blockwise_scaled_mm.argtypes = [ctypes.c_void_p] * 4 + [ctypes.c_int64] * 7 + [ctypes.c_void_p]
        L.oracle_check_division_scheme.restype = ctypes.c_int64
        L.oracle_check_division_scheme.argtypes = [ctypes.c_uint32, ctypes.c_uint32, ctypes.c_float,
                                                   ctypes.c_float, ctypes.c_int, ctypes.POINTER(ctypes.c_int64)]
        _lib = L
    return _lib


def _p(a: np.ndarray):
    return a.ctypes.data_as(ctypes.c_void_p)


def per_token_group_quant(x_bits: np.ndarray, in_dtype: str, group: int, eps: float = 1e-10,
                          dst_int8: bool = False, ue8m0: bool = False):
    """x_bits: [M,K] uint16 (bf16/f16 bit patterns) or float32.  Returns (codes u8 [M,K], scales f32 [M,G], sbyte u8)."""
    x_bits = np.ascontiguousarray(x_bits)
    M, K = x_bits.shape
    q = np.empty((M, K), np.uint8)
    s = np.empty((M, K // group), np.float32)
    sb = np.zeros((M, K // group), np.uint8)
    lib().oracle_per_token_group_quant(_p(x_bits), DT[in_dtype], M, K, group, eps, int(dst_int8),
                                       int(ue8m0), _p(q), _p(s), _p(sb))
    return q, s, sb


def blockwise_scaled_mm(a: np.ndarray, bt: np.ndarray, sa: np.ndarray, sb: np.ndarray) -> np.ndarray:
    """a [M,K] u8 codes, bt [N,K] u8 codes, sa [M,G] f32 (any strides), sb [G,NB] f32 (any strides)."""
    a = np.ascontiguousarray(a)
    bt = np.ascontiguousarray(bt)
    M, K = a.shape
    N = bt.shape[0]
    out = np.empty((M, N), np.float32)
    lib().oracle_blockwise_scaled_mm(_p(a), _p(bt), _p(sa), _p(sb), M, N, K,
                                     sa.strides[0] // 4, sa.strides[1] // 4,
                                     sb.strides[0] // 4, sb.strides[1] // 4, _p(out))
    return out


def check_division_scheme(amax_lo: int, amax_hi: int, qmax: float = 448.0, eps: float = 1e-10, src_f16: bool = False):
    n = ctypes.c_int64(0)
    bad = lib().oracle_check_division_scheme(amax_lo, amax_hi, qmax, eps, int(src_f16), ctypes.byref(n))
    return int(bad), int(n.value)
