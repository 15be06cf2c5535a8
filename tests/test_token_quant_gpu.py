"""GPU parity tests for sgl_per_token_quant_fp8 (torch.ops -> C ABI -> CUDA kernel).
Bar: BIT-EXACT codes and scales against the CPU oracle, on the committed golden vectors (made by the reference's
own torch check, tests/test_per_token_quant_fp8.py:17-25) and on the reference test's grid (:44-63: tokens
{128, 256, 512, 8192} x hidden {512, 1076, 1368, 2048, 4096}, torch.rand, seed 42, f16 and bf16)."""
import itertools

import pytest
import torch

from conftest import from_bits
from oracle import torch_oracle as O

import sgl_kernel

pytestmark = pytest.mark.gpu
DEV = "cuda"


def run(x_cpu):
    x = x_cpu.to(DEV)
    # the reference test's own allocation (tests/test_per_token_quant_fp8.py:31-35)
    scale = torch.zeros(x.size(0), device=DEV, dtype=torch.float32)
    out = torch.empty_like(x, dtype=torch.float8_e4m3fn)
    sgl_kernel.sgl_per_token_quant_fp8(x, out, scale)
    torch.cuda.synchronize()
    return out.cpu(), scale.cpu()


def assert_exact(q, s, q_ref, s_ref):
    nbad = int((s.view(torch.int32) != s_ref.view(torch.int32)).sum())
    assert nbad == 0, f"{nbad} / {s.numel()} scales differ"
    nbad = int((q.view(torch.uint8) != q_ref.view(torch.uint8)).sum())
    assert nbad == 0, f"{nbad} / {q.numel()} codes differ"


GOLDEN = [("ptq_128x512_bf16_rand", torch.bfloat16), ("ptq_16x1076_bf16_rand", torch.bfloat16),
          ("ptq_8x1368_f16_rand", torch.float16), ("ptq_4x7168_bf16_randn", torch.bfloat16),
          ("ptq_64x2048_f16_randn", torch.float16), ("ptq_3x20480_bf16_randn", torch.bfloat16)]


@pytest.mark.parametrize("name,sdt", GOLDEN)
def test_golden(token_golden, name, sdt):
    x = from_bits(token_golden[name + ".x"], sdt)
    q, s = run(x)
    assert_exact(q, s, torch.from_numpy(token_golden[name + ".q"]), torch.from_numpy(token_golden[name + ".s"]))


@pytest.mark.parametrize("dtype", [torch.float16, torch.bfloat16])
@pytest.mark.parametrize("tokens,hidden", list(itertools.product([128, 256, 512, 8192], [512, 1076, 1368, 2048, 4096])))
def test_reference_grid(dtype, tokens, hidden):
    torch.manual_seed(42)
    x = torch.rand((tokens, hidden), dtype=dtype)
    q, s = run(x)
    q_ref, s_ref = O.per_token_quant_fp8_ref(x)
    assert_exact(q, s, q_ref, s_ref)
    # and the reference's own assertion (:60-62)
    torch.testing.assert_close(q.float(), q_ref.float(), rtol=1e-3, atol=1e-3)


@pytest.mark.parametrize("tokens,hidden", [(1, 4), (1, 7168), (3, 8), (7, 12), (5, 16384), (2, 16392), (3, 40960),
                                           (2368, 2048), (2369, 1028), (4096, 7168), (33, 2052)])
def test_shapes_signed(tokens, hidden):
    """every dispatch arm: warp-per-token, CTA-per-token with 1/2/4/8 vectors per thread, two-pass rows, 8-byte vectors."""
    torch.manual_seed(tokens * 131 + hidden)
    x = torch.randn((tokens, hidden), dtype=torch.bfloat16) * 3.0
    q, s = run(x)
    q_ref, s_ref = O.per_token_quant_fp8_ref(x)
    assert_exact(q, s, q_ref, s_ref)


def test_zero_rows_negative_zero_and_extremes():
    x = torch.randn(8, 512, dtype=torch.bfloat16)
    x[0] = 0.0                      # scale 0 -> codes 0 (the kernel's guard, per_token_quant_fp8.cpp:110)
    x[1] = -0.0
    x[2, 5] = -0.0
    x[3] *= 1e-30
    x[4] *= 1e30
    x[5] = 0.0
    x[5, 7] = -2.0                  # one-hot row: that element is exactly -448
    q, s = run(x)
    q_ref, s_ref = O.per_token_quant_fp8_ref(x)
    assert_exact(q, s, q_ref, s_ref)
    assert q[5, 7].float() == -448.0 and s[0] == 0


def test_misaligned_views_use_the_narrow_path():
    # a row-sliced contiguous view whose base is only 8-byte aligned
    torch.manual_seed(1)
    big = torch.randn(9 * 520 + 4, dtype=torch.float16)
    x = big[4:].view(9, 520)
    xd = big.to(DEV)[4:].view(9, 520)
    assert xd.data_ptr() % 16 == 8
    scale = torch.zeros(9, device=DEV)
    out = torch.empty((9, 520), device=DEV, dtype=torch.float8_e4m3fn)
    sgl_kernel.sgl_per_token_quant_fp8(xd, out, scale)
    q_ref, s_ref = O.per_token_quant_fp8_ref(x.contiguous())
    assert_exact(out.cpu(), scale.cpu(), q_ref, s_ref)


def test_errors_before_launch():
    x = torch.randn(4, 64, dtype=torch.bfloat16, device=DEV)
    q = torch.empty_like(x, dtype=torch.float8_e4m3fn)
    s = torch.zeros(4, device=DEV)
    with pytest.raises(RuntimeError, match="BFloat16/Half"):
        sgl_kernel.sgl_per_token_quant_fp8(x.float(), q, s)
    with pytest.raises(RuntimeError, match="2D"):
        sgl_kernel.sgl_per_token_quant_fp8(x.view(2, 2, 64), q, s)
    with pytest.raises(RuntimeError, match="numel == num_tokens"):
        sgl_kernel.sgl_per_token_quant_fp8(x, q, torch.zeros(5, device=DEV))
    with pytest.raises(RuntimeError, match="divisible by 4"):
        sgl_kernel.sgl_per_token_quant_fp8(x[:, :62].contiguous(), q[:, :62].contiguous(), s)
    with pytest.raises(RuntimeError, match="Float8_e4m3fn"):
        sgl_kernel.sgl_per_token_quant_fp8(x, q.view(torch.uint8), s)
    # empty input is a no-op (per_token_quant_fp8.cpp:223)
    sgl_kernel.sgl_per_token_quant_fp8(x[:0], q[:0], s[:0])


def test_full_size_row_sample():
    """BASELINE-size activations (32768 x 7168): exact on a strided row sample."""
    torch.manual_seed(0)
    x = torch.randn((32768, 7168), dtype=torch.bfloat16, device=DEV)
    scale = torch.zeros(32768, device=DEV)
    out = torch.empty_like(x, dtype=torch.float8_e4m3fn)
    sgl_kernel.sgl_per_token_quant_fp8(x, out, scale)
    rows = torch.arange(0, 32768, 257, device=DEV)
    q_ref, s_ref = O.per_token_quant_fp8_ref(x[rows].cpu())
    assert_exact(out[rows].cpu(), scale[rows].cpu(), q_ref, s_ref)
