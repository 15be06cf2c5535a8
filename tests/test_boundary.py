"""CPU suite, part 2: the drop-in boundary.  The C-ABI library loads and exports every symbol that
include/sglk_fp8.h declares, the op schemas are the reference's verbatim, the python signatures match
the reference's, and argument validation fails loudly *before* any launch (no GPU is touched here)."""
import ctypes
import inspect
import os
import re

import pytest
import torch

import sgl_kernel
from sgl_kernel import _lib, _ops

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared_symbols():
    hdr = open(os.path.join(ROOT, "include", "sglk_fp8.h")).read()
    return re.findall(r"SGLK_API\s+[\w\s\*]+?\b(sglk_\w+)\s*\(", hdr)


def _exported_symbols(path):
    import subprocess

    out = subprocess.run(["nm", "-D", "--defined-only", path], capture_output=True, text=True, check=True).stdout
    return {ln.split()[-1] for ln in out.splitlines() if " T " in ln}


def test_library_exports_exactly_the_declared_symbols():
    """header == exports == python binding table, in both directions: nothing undeclared is reachable through the
    product library (laboratory code lives in libsglk_diag.so, built only by `make diag`)."""
    names = _declared_symbols()
    assert len(names) >= 10
    lib = ctypes.CDLL(_lib.LIB_PATH)
    for n in names:
        assert hasattr(lib, n), f"{n} declared in include/sglk_fp8.h but not exported"
    assert set(names) == set(_lib.SYMBOLS), "python binding table and header disagree"
    exported = {n for n in _exported_symbols(_lib.LIB_PATH) if not n.startswith("_")}
    assert exported == set(names), f"exported but undeclared: {sorted(exported - set(names))}; missing: {sorted(set(names) - exported)}"
    assert not any("debug" in n or "diag" in n for n in exported)
    assert _lib.load().sglk_abi_version() == _lib.ABI_VERSION


def test_error_strings_and_variants():
    lib = _lib.load()
    assert lib.sglk_error_string(0) == b"ok"
    for code in range(-9, 0):
        assert len(lib.sglk_error_string(code)) > 3
    n = lib.sglk_gemm_num_variants()
    assert n >= 4 and all(lib.sglk_gemm_variant_name(i) for i in range(n))
    assert lib.sglk_gemm_variant_name(n) == b""
    assert 0 <= lib.sglk_gemm_pick_variant(16384, 18432, 7168) < n
    assert 0 <= lib.sglk_gemm_pick_variant(1, 2112, 7168) < n


def test_c_abi_rejects_bad_arguments_before_launch():
    """Negative codes come back without any CUDA call, so this runs on a GPU-less box."""
    lib = _lib.load()
    buf = (ctypes.c_uint8 * 4096)()
    p = ctypes.addressof(buf)
    p = (p + 15) & ~15
    q = lib.sglk_per_token_group_quant_8bit
    ok = dict(num_groups=1, gpr=1, gs=128, idt=_lib.BF16, odt=_lib.E4M3)
    assert q(None, p, p, 1, 1, 128, _lib.BF16, _lib.E4M3, 1e-10, -448.0, 448.0, 0, 0, 0, None) == -1      # null
    assert q(p, p, p, 1, 1, 128, 7, _lib.E4M3, 1e-10, -448.0, 448.0, 0, 0, 0, None) == -2                 # in dtype
    assert q(p, p, p, 1, 1, 128, _lib.BF16, _lib.BF16, 1e-10, -448.0, 448.0, 0, 0, 0, None) == -2         # out dtype
    assert q(p, p, p, 1, 1, 100, _lib.BF16, _lib.E4M3, 1e-10, -448.0, 448.0, 0, 0, 0, None) == -3         # group
    assert q(p, p, p, 3, 2, 128, _lib.BF16, _lib.E4M3, 1e-10, -448.0, 448.0, 0, 0, 0, None) == -4         # groups % row
    assert q(p, p, p, 1, 1, 128, _lib.BF16, _lib.E4M3, 1e-10, -448.0, 448.0, 0, 9, 0, None) == -6         # layout
    assert q(p, p, p, 1, 1, 128, _lib.BF16, _lib.E4M3, 1e-10, -448.0, 448.0, 1, 0, 0, None) == -6         # ue8m0 vs layout
    assert q(p + 1, p, p, 1, 1, 128, _lib.BF16, _lib.E4M3, 1e-10, -448.0, 448.0, 0, 0, 0, None) == -5     # element alignment
    assert q(p, p, p, 0, 1, 128, _lib.BF16, _lib.E4M3, 1e-10, -448.0, 448.0, 0, 0, 0, None) == 0          # empty: no-op
    g = lib.sglk_fp8_blockwise_scaled_mm
    assert g(p, p, p, p, p, 4, 128, 100, 112, 112, 128, 1, 4, 1, 1, _lib.BF16, None) == -5                # K % 16
    assert g(p, p, p, p, p, 4, 128, 128, 128, 128, 128, 1, 4, 1, 1, _lib.F32, None) == -2                 # out dtype
    assert g(p, p, p, p, p, -1, 128, 128, 128, 128, 128, 1, 4, 1, 1, _lib.BF16, None) == -4               # shape
    assert g(None, p, p, p, p, 4, 128, 128, 128, 128, 128, 1, 4, 1, 1, _lib.BF16, None) == -1             # null
    assert g(p, p, p, p, p, 0, 128, 128, 128, 128, 128, 1, 4, 1, 1, _lib.BF16, None) == 0                 # empty
    ex = lib.sglk_fp8_blockwise_scaled_mm_ex
    assert ex(p, p, p, p, p, 4, 128, 128, 128, 128, 128, 1, 4, 1, 1, _lib.BF16, 99, 0, 0, None, 0, None) == -9     # variant
    for v in range(lib.sglk_gemm_num_variants()):       # laboratory ids (diagnostics with wrong results) are not launchable
        if not lib.sglk_gemm_variant_is_product(v):
            assert ex(p, p, p, p, p, 4, 128, 128, 128, 128, 128, 1, 4, 1, 1, _lib.BF16, v, 0, 0, None, 0, None) == -9
            assert lib.sglk_gemm_variant_name(v).startswith((b"diag_", b"lab_"))


def test_schemas_are_the_reference_strings():
    # /root/reference/src/torch_extension_sycl.cc:395-402 (verbatim, whitespace-normalised)
    want_v1 = ("sgl_per_token_group_quant_8bit(Tensor input, Tensor output_q, Tensor output_s, int group_size,"
               " float eps, float fp8_min, float fp8_max, bool scale_ue8m0) -> ()")
    want_v2 = ("sgl_per_token_group_quant_8bit_v2(Tensor input, Tensor output_q, Tensor output_s, int group_size,"
               " float eps, float fp8_min, float fp8_max, bool scale_ue8m0, bool fuse_silu_and_mul, Tensor? masked_m) -> ()")
    assert _ops._QUANT_SCHEMA == want_v1 and _ops._QUANT_V2_SCHEMA == want_v2
    s1 = torch.ops.sgl_kernel.sgl_per_token_group_quant_8bit.default._schema
    assert [a.name for a in s1.arguments] == ["input", "output_q", "output_s", "group_size", "eps", "fp8_min",
                                              "fp8_max", "scale_ue8m0"]
    assert len(s1.returns) == 0
    s2 = torch.ops.sgl_kernel.sgl_per_token_group_quant_8bit_v2.default._schema
    assert [a.name for a in s2.arguments][-2:] == ["fuse_silu_and_mul", "masked_m"]
    s3 = torch.ops.sgl_kernel.fp8_blockwise_scaled_mm.default._schema
    assert [a.name for a in s3.arguments] == ["mat_a", "mat_b", "scales_a", "scales_b", "out_dtype"]
    assert "ScalarType out_dtype" in str(s3) and len(s3.returns) == 1


def test_python_signatures_match_reference():
    # python/sgl_kernel/gemm.py:24 and :85-97
    assert list(inspect.signature(sgl_kernel.fp8_blockwise_scaled_mm).parameters) == [
        "mat_a", "mat_b", "scales_a", "scales_b", "out_dtype"]
    sig = inspect.signature(sgl_kernel.sgl_per_token_group_quant_8bit)
    assert list(sig.parameters) == ["input", "output_q", "output_s", "group_size", "eps", "fp8_min", "fp8_max",
                                    "scale_ue8m0", "fuse_silu_and_mul", "masked_m", "enable_v2"]
    assert sig.parameters["scale_ue8m0"].default is False and sig.parameters["enable_v2"].default is None
    assert sgl_kernel.sgl_per_token_group_quant_fp8 is sgl_kernel.sgl_per_token_group_quant_8bit
    assert sgl_kernel.sgl_per_token_group_quant_int8 is sgl_kernel.sgl_per_token_group_quant_8bit


def test_no_cpu_fallback():
    """CPU tensors must not be silently computed: the ops exist for dispatch key CUDA only."""
    x = torch.randn(4, 256, dtype=torch.bfloat16)
    q = torch.empty(4, 256, dtype=torch.float8_e4m3fn)
    s = torch.empty(4, 2)
    with pytest.raises((NotImplementedError, RuntimeError)):
        sgl_kernel.sgl_per_token_group_quant_8bit(x, q, s, 128, 1e-10, -448.0, 448.0, enable_v2=False)
    with pytest.raises((NotImplementedError, RuntimeError)):
        sgl_kernel.fp8_blockwise_scaled_mm(q, q.t(), s, s.t(), torch.bfloat16)


def test_v1_asserts_like_reference():
    x = torch.empty(0)
    with pytest.raises(AssertionError, match="only v2 support fuse_silu_and_mul"):
        sgl_kernel.sgl_per_token_group_quant_8bit(x, x, x, 128, 1e-10, -448.0, 448.0, False, True, None, False)
    with pytest.raises(AssertionError, match="only v2 support masked_m"):
        sgl_kernel.sgl_per_token_group_quant_8bit(x, x, x, 128, 1e-10, -448.0, 448.0, False, False, x, False)


def test_env_switch_selects_v2(monkeypatch):
    calls = []
    monkeypatch.setattr(torch.ops.sgl_kernel.sgl_per_token_group_quant_8bit_v2, "default", lambda *a: calls.append("v2"))
    monkeypatch.setattr(torch.ops.sgl_kernel.sgl_per_token_group_quant_8bit, "default", lambda *a: calls.append("v1"))
    x = torch.empty(0)
    monkeypatch.setenv("SGLANG_PER_TOKEN_GROUP_QUANT_8BIT_V2", "1")
    sgl_kernel.sgl_per_token_group_quant_8bit(x, x, x, 128, 1e-10, -448.0, 448.0)
    monkeypatch.setenv("SGLANG_PER_TOKEN_GROUP_QUANT_8BIT_V2", "false")
    sgl_kernel.sgl_per_token_group_quant_8bit(x, x, x, 128, 1e-10, -448.0, 448.0)
    assert calls == ["v2", "v1"]


def test_scale_allocation_layouts():
    f = sgl_kernel.create_per_token_group_quant_fp8_output_scale
    s = f((5, 7168), "cpu", 128, False, False, False)
    assert s.shape == (5, 56) and s.is_contiguous()
    s = f((5, 7168), "cpu", 128, True, False, False)
    assert s.shape == (5, 56) and s.stride() == (1, 5)
    s = f((5, 7168), "cpu", 128, True, True, False)
    assert s.shape == (5, 56) and s.stride() == (1, 8)
    s = f((5, 7168), "cpu", 128, True, True, True)
    assert s.shape == (5, 14) and s.dtype == torch.int32 and s.stride() == (1, 8)


def test_meta_kernels_trace():
    a = torch.empty(8, 256, dtype=torch.float8_e4m3fn, device="meta")
    b = torch.empty(128, 256, dtype=torch.float8_e4m3fn, device="meta").t()
    sa = torch.empty(8, 2, device="meta")
    sb = torch.empty(2, 1, device="meta")
    out = sgl_kernel.fp8_blockwise_scaled_mm(a, b, sa, sb, torch.bfloat16)
    assert out.shape == (8, 128) and out.dtype == torch.bfloat16 and out.device.type == "meta"


def test_product_never_imports_oracle():
    """The product path must not route through the CPU oracle (or /root/reference)."""
    pkg = os.path.join(ROOT, "sgl-kernel-xpu_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                txt = open(os.path.join(dirpath, f)).read()
                assert not re.search(r"^\s*(from|import)\s+oracle\b", txt, re.M), f"{f} imports the oracle"
                assert "torch_oracle" not in txt and "c_oracle" not in txt, f


def test_grouped_bench_imports_under_a_foreign_command_line(monkeypatch):
    """bench.py imports tools/grouped_bench.py for its f4 sub-record; the module must not read the importer's argv
    (BENCH_r02.json carried "invalid literal for int() with base 10: '--gpus'" instead of the record) nor touch a GPU
    at import time."""
    import importlib
    import sys

    monkeypatch.setattr(sys, "argv", ["bench.py", "--gpus", "1", "--steps", "20"])
    monkeypatch.syspath_prepend(os.path.join(ROOT, "tools"))
    sys.modules.pop("grouped_bench", None)
    gb = importlib.import_module("grouped_bench")
    assert gb.Es == [8, 32, 64] and gb.avg == 512 and callable(gb.records)
    sys.modules.pop("grouped_bench", None)


def test_jit_entry_points_have_the_reference_signatures():
    """sgl_kernel.jit.per_token_group_quant_8bit{,_v2}: names, order and defaults of
    /root/reference/python/sgl_kernel/jit/per_token_group_quant_8bit.py:153-162 and
    jit/per_token_group_quant_8bit_v2.py:118-129; both refuse a CPU tensor the way the reference refuses a non-XPU one."""
    from sgl_kernel import jit

    assert set(jit.__all__) == {"per_token_group_quant_8bit", "per_token_group_quant_8bit_v2"}
    common = ["input", "output_q", "output_s", "group_size", "eps", "min_8bit", "max_8bit", "scale_ue8m0"]
    s1 = inspect.signature(jit.per_token_group_quant_8bit)
    assert list(s1.parameters) == common and s1.parameters["scale_ue8m0"].default is False
    s2 = inspect.signature(jit.per_token_group_quant_8bit_v2)
    assert list(s2.parameters) == common + ["fuse_silu_and_mul", "masked_m"]
    assert s2.parameters["fuse_silu_and_mul"].default is False and s2.parameters["masked_m"].default is None
    x = torch.zeros(2, 128, dtype=torch.bfloat16)
    q, s = torch.empty(2, 128, dtype=torch.float8_e4m3fn), torch.empty(2, 1)
    with pytest.raises(RuntimeError, match="CUDA"):
        jit.per_token_group_quant_8bit(x, q, s, 128, 1e-10, -448.0, 448.0)
    with pytest.raises(RuntimeError, match="CUDA"):
        jit.per_token_group_quant_8bit_v2(x, q, s, 128, 1e-10, -448.0, 448.0)
    with pytest.raises(AssertionError, match="eps"):
        jit.per_token_group_quant_8bit_v2(x, q, s, 128, 1e-6, -448.0, 448.0)
