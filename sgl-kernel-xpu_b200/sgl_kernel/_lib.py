"""ctypes loader for libsglk_sm100a.so — the C ABI declared in include/sglk_fp8.h.

Mirrors the reference's JIT calling convention (python/sgl_kernel/jit/per_token_group_quant_8bit.py:77-150:
a cached ctypes handle whose functions take raw data pointers, scalars and the stream handle).
There is deliberately no fallback: if the library is missing or fails to load, every op raises.
"""
from __future__ import annotations

import ctypes
import os

_PKG_DIR = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get(
    "SGLK_LIBRARY", os.path.join(os.path.dirname(_PKG_DIR), "lib", "libsglk_sm100a.so")
)

# dtype / layout codes of include/sglk_fp8.h
BF16, F16, F32, E4M3, INT8 = 0, 1, 2, 3, 4
SCALE_F32_ROW_MAJOR, SCALE_F32_COL_MAJOR, SCALE_UE8M0_ROW_MAJOR, SCALE_UE8M0_COL_PACKED = 0, 1, 2, 3
ABI_VERSION = 3
STREAMK_AUTO, STREAMK_FORCE, STREAMK_NEVER, GROUP_M_SHIFT, SPLIT_SHIFT, DECODE_2D_LOADS = 0, 1, 2, 8, 16, 1 << 24

_i64, _i32, _f32, _vp = ctypes.c_int64, ctypes.c_int, ctypes.c_float, ctypes.c_void_p

# name -> (restype, argtypes): every symbol include/sglk_fp8.h declares
SYMBOLS = {
    "sglk_abi_version": (_i32, []),
    "sglk_error_string": (ctypes.c_char_p, [_i32]),
    "sglk_per_token_group_quant_8bit": (
        _i32, [_vp, _vp, _vp, _i64, _i64, _i64, _i32, _i32, _f32, _f32, _f32, _i32, _i32, _i64, _vp]),
    "sglk_silu_mul_per_token_group_quant_8bit": (
        _i32, [_vp, _vp, _vp, _i64, _i64, _i64, _i32, _i32, _f32, _f32, _f32, _i32, _i32, _i64, _vp]),
    "sglk_per_token_group_quant_8bit_v2": (
        _i32, [_vp, _vp, _vp, _i64, _i64, _i64, _i64, _i32, _i32, _f32, _f32, _f32, _i32, _i32, _i64, _i64, _i32, _vp, _vp]),
    "sglk_per_token_quant_fp8": (_i32, [_vp, _vp, _vp, _i64, _i64, _i32, _vp]),
    "sglk_fp8_blockwise_scaled_mm": (
        _i32, [_vp, _vp, _vp, _vp, _vp] + [_i64] * 10 + [_i32, _vp]),
    "sglk_fp8_blockwise_scaled_mm_ex": (
        _i32, [_vp, _vp, _vp, _vp, _vp] + [_i64] * 10 + [_i32, _i32, _i32, _i32, _vp, _i64, _vp]),
    "sglk_fp8_blockwise_scaled_mm_quant": (
        _i32, [_vp, _vp, _vp, _vp, _vp, _vp] + [_i64] * 12 + [_i32, _f32, _i32, _i32, _i32, _vp, _i64, _vp]),
    "sglk_fp8_blockwise_scaled_grouped_mm": (
        _i32, [_vp, _vp, _vp, _vp, _vp, _vp, _vp] + [_i64] * 12 + [_i32, _i32, _vp, _i64, _vp]),
    "sglk_grouped_mm_workspace_bytes": (_i64, [_i64, _i64]),
    "sglk_fp8_blockwise_scaled_grouped_mm_static": (
        _i32, [_vp, _vp, _vp, _vp, _vp, _vp, _vp] + [_i64] * 12 + [_i32, _i32, _vp]),
    "sglk_grouped_mm_static_max_tiles": (_i64, []),
    "sglk_gemm_workspace_bytes": (_i64, [_i64, _i64, _i64, _i32, _i32, _i32]),
    "sglk_fp8_blockwise_scaled_mm_allgather": (
        _i32, [_vp, _vp, _vp, _vp, _vp, ctypes.POINTER(_vp), _i32, _vp] + [_i64] * 10 + [_i32, _i32, _i32, _vp, _i64, _vp]),
    "sglk_per_token_group_quant_fp8_bcast": (_i32, [_vp, _vp, _vp, _vp, _i64, _i64, _i64, _i32, _f32, _i32, _vp]),
    "sglk_fp8_blockwise_scaled_mm_gated": (
        _i32, [_vp, _vp, _vp, _vp, _vp] + [_i64] * 10 + [_i32, _i32, _vp, ctypes.c_uint32, _i32, _i32, _i32, _i64, _vp]),
    "sglk_gated_walk_order": (_i64, [_vp, _i64, _i64, _i64, _i32, _i32, _i32]),
    "sglk_gemm_num_variants": (_i32, []),
    "sglk_gemm_variant_name": (ctypes.c_char_p, [_i32]),
    "sglk_gemm_variant_is_product": (_i32, [_i32]),
    "sglk_gemm_pick_variant": (_i32, [_i64, _i64, _i64]),
    "sglk_selftest_division": (_i32, [_vp, _i64, _i32, _vp, _vp]),
}

_lib = None


def load() -> ctypes.CDLL:
    """Load (once) and type the library.  Raises RuntimeError if it is not built."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise RuntimeError(
            f"sgl_kernel (B200): {LIB_PATH} not found — build it with `python -c 'import __graft_entry__ as g; "
            f"g.build()'` or `make -C sgl-kernel-xpu_b200/csrc`.  There is no CPU or PyTorch fallback."
        )
    lib = ctypes.CDLL(LIB_PATH)
    for name, (res, args) in SYMBOLS.items():
        fn = getattr(lib, name)  # AttributeError here == ABI mismatch: fail loudly
        fn.restype = res
        fn.argtypes = args
    if lib.sglk_abi_version() != ABI_VERSION:
        raise RuntimeError(f"sgl_kernel (B200): ABI version mismatch in {LIB_PATH}")
    _lib = lib
    return lib


def check(code: int, what: str) -> None:
    """Turn a non-zero return of the C ABI into the RuntimeError a TORCH_CHECK would raise."""
    if code != 0:
        msg = load().sglk_error_string(code).decode()
        raise RuntimeError(f"{what}: {msg} (code {code})")
