#!/usr/bin/env python
"""Generate the golden vectors under tests/golden/ by RUNNING THE REFERENCE'S OWN CODE.

The reference's test modules import `sgl_kernel` / `torch.xpu` at module top and cannot be
imported, so the pure-torch oracle functions are lifted out of them by AST (no text is copied
into this repo) and executed here, on CPU, on seeded inputs.  Inputs and outputs are stored as
raw bit patterns in small .npz files; `tests/test_oracle.py` pins `oracle/` against them and the
`-m gpu` tests pin the CUDA path against them.

Run in the build container only (needs /root/reference):   python tests/golden/make_golden.py
"""
import ast
import os
import sys
from typing import List, Optional, Tuple, Type  # noqa: F401  (names used by the lifted code)

import numpy as np
import torch

REF = os.environ.get("SGLK_REFERENCE", "/root/reference")
OUT = os.path.dirname(os.path.abspath(__file__))


def lift(path: str, names):
    """Compile the named top-level functions of a reference file into a fresh namespace."""
    src = open(os.path.join(REF, path)).read()
    tree = ast.parse(src)
    keep = [n for n in tree.body if isinstance(n, ast.FunctionDef) and n.name in names]
    missing = set(names) - {n.name for n in keep}
    assert not missing, f"{path}: missing {missing}"
    ns = {"torch": torch, "Tuple": Tuple, "Optional": Optional, "Type": Type, "List": List}
    exec(compile(ast.Module(body=keep, type_ignores=[]), path, "exec"), ns)
    return ns


def bits(t: torch.Tensor) -> np.ndarray:
    t = t.contiguous()
    if t.dtype in (torch.bfloat16, torch.float16):
        return t.view(torch.int16).numpy().view(np.uint16)
    if t.dtype == torch.float8_e4m3fn:
        return t.view(torch.uint8).numpy()
    return t.numpy()


def main():
    q = lift(
        "tests/test_per_token_group_quant_8bit.py",
        ["ceil_to_ue8m0", "per_token_group_quant_fp8_ref", "per_token_group_quant_int8_ref"],
    )
    g = lift("tests/test_fp8_blockwise_gemm.py", ["cdiv", "scale_shape", "baseline_scaled_mm"])
    w = lift("tests/test_es_fp8_blockwise_moe.py", ["ceil_div", "per_block_cast_to_fp8"])
    v2 = lift("tests/test_per_token_group_quant_8bit_v2.py", ["per_token_group_quant_fp8_ref"])

    # ---------------- quant: shapes / seeds of tests/test_per_token_group_quant_8bit.py:278-355
    quant = {}
    cases = [
        # name, M, K, group, src dtype, dst, seed, scale factor
        ("fp8_bf16_128x1024_g128", 128, 1024, 128, torch.bfloat16, "fp8", 42, 1.0),
        ("fp8_bf16_64x512_g64", 64, 512, 64, torch.bfloat16, "fp8", 42, 1.0),
        ("fp8_bf16_5x7168_g128", 5, 7168, 128, torch.bfloat16, "fp8", 42, 1.0),
        ("fp8_bf16_40x2048_g128", 40, 2048, 128, torch.bfloat16, "fp8", 42, 1.0),
        ("fp8_bf16_32x256_g32", 32, 256, 32, torch.bfloat16, "fp8", 0, 1.0),
        ("fp8_bf16_3x1024_g256", 3, 1024, 256, torch.bfloat16, "fp8", 7, 1.0),
        ("fp8_bf16_2x1024_g512", 2, 1024, 512, torch.bfloat16, "fp8", 7, 1.0),
        ("fp8_f16_128x1024_g128", 128, 1024, 128, torch.float16, "fp8", 42, 1.0),
        ("fp8_f32_64x512_g64", 64, 512, 64, torch.float32, "fp8", 42, 1.0),
        ("fp8_bf16_64x512_g64_small", 64, 512, 64, torch.bfloat16, "fp8", 42, 1e-3),
        ("fp8_bf16_64x512_g64_large", 64, 512, 64, torch.bfloat16, "fp8", 42, 100.0),
        ("int8_bf16_128x1024_g128", 128, 1024, 128, torch.bfloat16, "int8", 42, 1.0),
        ("int8_f16_128x1024_g128", 128, 1024, 128, torch.float16, "int8", 42, 1.0),
        ("int8_bf16_16x4096_g64", 16, 4096, 64, torch.bfloat16, "int8", 42, 1.0),
    ]
    for name, M, K, gs, sdt, dst, seed, sf in cases:
        torch.manual_seed(seed)
        x = torch.randn(M, K, dtype=sdt)
        if sf != 1.0:
            x = x * sf
        if dst == "fp8":
            xq, xs = q["per_token_group_quant_fp8_ref"](x, gs, 1e-10, False)
        else:
            xq, xs = q["per_token_group_quant_int8_ref"](x, gs, 1e-10)
        quant[name + ".x"] = bits(x)
        quant[name + ".q"] = bits(xq)
        quant[name + ".s"] = bits(xs)

    # eps path / mixed magnitudes / negative zero: groups of zeros, tiny, huge, one-hot
    torch.manual_seed(3)
    x = torch.randn(8, 512, dtype=torch.bfloat16)
    x[0, :128] = 0.0
    x[1, 128:256] = -0.0
    x[2, :128] *= 1e-30
    x[3, :128] *= 1e30
    x[4, :128] = 0.0
    x[4, 5] = 1.0
    x[5, :] = x[5, :].abs() * 1e-12          # below eps*448 after division
    x[6, 128:256] = 2.0 ** -126               # smallest fp32 normal representable in bf16
    xq, xs = q["per_token_group_quant_fp8_ref"](x, 128, 1e-10, False)
    quant["fp8_bf16_edge_8x512_g128.x"] = bits(x)
    quant["fp8_bf16_edge_8x512_g128.q"] = bits(xq)
    quant["fp8_bf16_edge_8x512_g128.s"] = bits(xs)

    # UE8M0: scales rounded up to powers of two (tests/..8bit.py:17-19,357-417, seed 0)
    for name, M, K, gs in [("ue8m0_128x1024_g128", 128, 1024, 128), ("ue8m0_5x2048_g128", 5, 2048, 128),
                           ("ue8m0_32x256_g32", 32, 256, 32)]:
        torch.manual_seed(0)
        x = torch.randn(M, K, dtype=torch.bfloat16)
        xq, xs = q["per_token_group_quant_fp8_ref"](x, gs, 1e-10, True)
        # byte formula asserted exactly by the reference test (:399-405)
        xv = x.view(M, K // gs, gs).float()
        amax = xv.abs().amax(dim=2).clamp(min=1e-10)
        exp_s = torch.ceil(torch.log2((amax / 448.0).clamp(min=1e-10)))
        byte_ref = (exp_s.to(torch.int32) + 127).to(torch.uint8)
        quant[name + ".x"] = bits(x)
        quant[name + ".q"] = bits(xq)
        quant[name + ".s"] = bits(xs)
        quant[name + ".sbyte"] = bits(byte_ref)

    # v2 fused silu*mul oracle (tests/test_per_token_group_quant_8bit_v2.py:408-444)
    torch.manual_seed(11)
    x = torch.randn(16, 2 * 512, dtype=torch.bfloat16)
    xq, xs = v2["per_token_group_quant_fp8_ref"](x, 128, 1e-10, False, True)
    quant["v2_silu_16x1024_g128.x"] = bits(x)
    quant["v2_silu_16x1024_g128.q"] = bits(xq)
    quant["v2_silu_16x1024_g128.s"] = bits(xs)
    np.savez_compressed(os.path.join(OUT, "quant_golden.npz"), **quant)

    # ---------------- GEMM: input recipe of tests/test_fp8_blockwise_gemm.py:66-80
    gemm = {}
    for M, N, K, odt, seed in [
        (1, 128, 512, torch.bfloat16, 0),
        (3, 256, 512, torch.float16, 1),
        (5, 512, 1024, torch.bfloat16, 2),
        (127, 128, 512, torch.bfloat16, 3),
        (128, 256, 1024, torch.float16, 4),
        (130, 384, 256, torch.bfloat16, 5),
    ]:
        torch.manual_seed(seed)
        fp8_max = 448.0
        a_fp32 = (torch.rand(M, K, dtype=torch.float32) - 0.5) * 2 * fp8_max
        a_fp8 = a_fp32.clamp(min=-fp8_max, max=fp8_max).to(torch.float8_e4m3fn)
        b_fp32 = (torch.rand(N, K, dtype=torch.float32) - 0.5) * 2 * fp8_max
        b_fp8 = b_fp32.clamp(min=-fp8_max, max=fp8_max).to(torch.float8_e4m3fn).t()
        sa_shape = g["scale_shape"](a_fp8.shape, (1, 128))
        sb_shape = g["scale_shape"](b_fp8.shape, (128, 128))
        scale_a = (torch.randn(sa_shape, dtype=torch.float32) * 0.001).t().contiguous().t()
        scale_b = (torch.randn(sb_shape, dtype=torch.float32) * 0.001).t().contiguous().t()
        o = g["baseline_scaled_mm"](a_fp8, b_fp8, scale_a, scale_b, odt)
        name = f"mm_{M}x{N}x{K}_{'bf16' if odt == torch.bfloat16 else 'f16'}"
        gemm[name + ".a"] = bits(a_fp8)
        gemm[name + ".bt"] = bits(b_fp8.t())             # stored as the contiguous [N,K] buffer
        gemm[name + ".sa"] = bits(scale_a)               # logical [M,G]
        gemm[name + ".sb"] = bits(scale_b)               # logical [G,N/128]
        gemm[name + ".o"] = bits(o)

    # realistic chain: randn weights -> per_block_cast_to_fp8; randn bf16 activations -> oracle quant
    torch.manual_seed(42)
    M, N, K = 8, 256, 512
    x = torch.randn(M, K, dtype=torch.bfloat16)
    wt = torch.randn(N, K)
    wq, ws = w["per_block_cast_to_fp8"](wt)
    xq, xs = q["per_token_group_quant_fp8_ref"](x, 128, 1e-10, False)
    o = g["baseline_scaled_mm"](xq, wq.t(), xs, ws.t(), torch.bfloat16)
    gemm["chain_8x256x512.x"] = bits(x)
    gemm["chain_8x256x512.w"] = bits(wt)
    gemm["chain_8x256x512.wq"] = bits(wq)
    gemm["chain_8x256x512.ws"] = bits(ws)
    gemm["chain_8x256x512.xq"] = bits(xq)
    gemm["chain_8x256x512.xs"] = bits(xs)
    gemm["chain_8x256x512.o"] = bits(o)
    np.savez_compressed(os.path.join(OUT, "gemm_golden.npz"), **gemm)

    for f in ("quant_golden.npz", "gemm_golden.npz"):
        print(f, os.path.getsize(os.path.join(OUT, f)), "bytes")
    print("torch", torch.__version__)


def main_grouped():
    """Grouped (MoE) blockwise GEMM: the recipe of tests/test_fp8_blockwise_moe.py:66-171 (its own `to_fp8` and
    `baseline_scaled_mm`, lifted), seeded and small; per-expert baselines are stored stacked."""
    import random

    m = lift("tests/test_fp8_blockwise_moe.py", ["cdiv", "scale_shape", "to_fp8", "baseline_scaled_mm"])
    out = {}
    for case, (E, n_g, k_g, odt, seed) in enumerate([(4, 256, 512, torch.bfloat16, 0), (8, 384, 256, torch.float16, 1),
                                                     (3, 128, 1024, torch.bfloat16, 2)]):
        random.seed(seed)
        torch.manual_seed(seed)
        offs, sizes, a_l, bt_l, sa_l, sbt_l, o_l = [0], [], [], [], [], [], []
        for g in range(E):
            m_g = 16 * random.randint(1, 20)
            offs.append(offs[-1] + m_g)
            sizes.append([m_g, n_g, k_g])
            a_g = m["to_fp8"](torch.randn((m_g, k_g)))
            b_g = m["to_fp8"](torch.randn((n_g, k_g)).t())
            sa_g = torch.randn(m["scale_shape"](a_g.shape, (1, 128))) * 0.001
            sb_g = torch.randn(m["scale_shape"](b_g.shape, (128, 128))) * 0.001
            o_l.append(m["baseline_scaled_mm"](a_g, b_g, sa_g, sb_g, odt))
            a_l.append(a_g); bt_l.append(b_g.t().contiguous()); sa_l.append(sa_g); sbt_l.append(sb_g.t().contiguous())
        name = f"grouped{case}_E{E}_n{n_g}_k{k_g}_{'bf16' if odt == torch.bfloat16 else 'f16'}"
        out[name + ".offsets"] = np.asarray(offs[:-1], dtype=np.int32)
        out[name + ".sizes"] = np.asarray(sizes, dtype=np.int32)
        out[name + ".a"] = bits(torch.cat(a_l))                    # [sum_m, K]
        out[name + ".bt"] = bits(torch.stack(bt_l))                # [E, N, K]
        out[name + ".sa"] = bits(torch.cat(sa_l))                  # [sum_m, K/128]
        out[name + ".sbt"] = bits(torch.stack(sbt_l))              # [E, N/128, K/128]
        out[name + ".o"] = bits(torch.cat(o_l))                    # [sum_m, N]
    np.savez_compressed(os.path.join(OUT, "grouped_golden.npz"), **out)
    print("grouped_golden.npz", os.path.getsize(os.path.join(OUT, "grouped_golden.npz")), "bytes")


def main_token():
    """sgl_per_token_quant_fp8: the reference test's own torch check (tests/test_per_token_quant_fp8.py:17-25, lifted)
    applied to the scales the kernel defines (rowmax|x| / 448, src/sycl/per_token_quant_fp8.cpp:105); inputs as the
    test draws them (torch.rand, seed 42, :52-53) plus signed / zero-row cases."""
    t = lift("tests/test_per_token_quant_fp8.py", ["torch_per_token_quant_fp8"])
    out = {}
    for tokens, hidden, dt, kind in [(128, 512, torch.bfloat16, "rand"), (16, 1076, torch.bfloat16, "rand"),
                                     (8, 1368, torch.float16, "rand"), (4, 7168, torch.bfloat16, "randn"),
                                     (64, 2048, torch.float16, "randn"), (3, 20480, torch.bfloat16, "randn")]:
        torch.manual_seed(42)
        x = (torch.rand if kind == "rand" else torch.randn)((tokens, hidden), dtype=dt)
        if kind == "randn":
            x[0, 1] = -0.0
        scale = x.float().abs().amax(dim=1) / 448.0
        q = t["torch_per_token_quant_fp8"](x, scale)
        name = f"ptq_{tokens}x{hidden}_{'bf16' if dt == torch.bfloat16 else 'f16'}_{kind}"
        out[name + ".x"] = bits(x)
        out[name + ".s"] = bits(scale)
        out[name + ".q"] = bits(q)
    np.savez_compressed(os.path.join(OUT, "token_golden.npz"), **out)
    print("token_golden.npz", os.path.getsize(os.path.join(OUT, "token_golden.npz")), "bytes")


if __name__ == "__main__":
    if "--token" in sys.argv:        # only the per-token fixture
        sys.exit(main_token())
    if "--grouped" in sys.argv:      # only the grouped fixture (the other two files are left as committed)
        sys.exit(main_grouped())
    main()
    main_grouped()
    sys.exit(main_token())
