"""The grouped GEMM for callers that cannot allocate: sglk_fp8_blockwise_scaled_grouped_mm_static.

Some users of this library are not allowed to create device memory when they call it — a plugin living inside another
program's memory pool, a server that has already given all of HBM to its caches, a cgo/JNI/ctypes binding that is handed
input and output pointers but owns no allocator.  The ordinary grouped GEMM wants a workspace from them for its m-tile
table; the `_static` entry point keeps that table in a 64 KiB array inside the library instead, so the caller passes only
inputs and outputs.

What is checked here, straight through the C ABI (ctypes, raw pointers — no torch op in between):
  * the `_static` call gives the same output as the workspace call (same kernels, so the bar is bit equality; the
    suite's usual grouped-GEMM tolerance rtol=1e-2 / atol=5e-4 is asserted as well);
  * on calls in a row that share the array: a problem with many tiles, then one with few (different tile shape, so the
    stale tail of the first table is still lying there), then the first again;
  * problems beyond the array's capacity are refused with SGLK_E_CAPACITY and a readable message, before any launch.

The file name sorts last on purpose: `pytest -x` stops at the first failure, and a problem in this add-on must not hide the
parity tests.  The GPU test passes both alone in a fresh process (its first call does the once-per-device address lookup)
and after the rest of the suite.
"""
import pytest
import torch

from sgl_kernel import _lib

DEV = "cuda"
E_CAPACITY = -10


def _problem(m_list, n, k, seed, gap=0):
    """Stacked expert activations (gap unused rows after each expert), weights and scales, all on the device."""
    g = torch.Generator(device=DEV).manual_seed(seed)
    E = len(m_list)
    rows = sum(m + gap for m in m_list)
    kb, nb = -(k // -128), -(n // -128)
    a = (torch.randn(rows, k, generator=g, device=DEV) * 40).to(torch.float8_e4m3fn)
    b = (torch.randn(E * n, k, generator=g, device=DEV) * 40).to(torch.float8_e4m3fn)      # expert e: rows [e*n, (e+1)*n), K contiguous
    sa = torch.randn(rows, kb, generator=g, device=DEV) * 1e-3
    sb = torch.randn(E, kb, nb, generator=g, device=DEV) * 1e-3
    offs, at = [], 0
    for m in m_list:
        offs.append(at)
        at += m + gap
    offsets = torch.tensor(offs, dtype=torch.int32, device=DEV)
    sizes = torch.tensor([[m, n, k] for m in m_list], dtype=torch.int32, device=DEV)
    return dict(a=a, b=b, sa=sa, sb=sb, offsets=offsets, sizes=sizes, E=E, rows=rows, n=n, k=k)


def _common_args(p, out):
    return [p["a"].data_ptr(), p["b"].data_ptr(), p["sa"].data_ptr(), p["sb"].data_ptr(), out.data_ptr(),
            p["offsets"].data_ptr(), p["sizes"].data_ptr(), p["E"], p["rows"], p["n"], p["k"],
            p["a"].stride(0), p["b"].stride(0), out.stride(0), p["sa"].stride(0), p["sa"].stride(1),
            p["sb"].stride(0), p["sb"].stride(1), p["sb"].stride(2), _lib.BF16, 0]


def _with_workspace(lib, p):
    out = torch.full((p["rows"], p["n"]), float("nan"), dtype=torch.bfloat16, device=DEV)
    need = lib.sglk_grouped_mm_workspace_bytes(p["E"], p["rows"])
    ws = torch.empty(need + 256, dtype=torch.uint8, device=DEV)
    ptr = (ws.data_ptr() + 15) & ~15
    stream = torch.cuda.current_stream().cuda_stream
    rc = lib.sglk_fp8_blockwise_scaled_grouped_mm(*_common_args(p, out), ptr, need, stream)
    assert rc == 0, lib.sglk_error_string(rc)
    return out


def _static(lib, p):
    out = torch.full((p["rows"], p["n"]), float("nan"), dtype=torch.bfloat16, device=DEV)   # the caller's only buffers: inputs + output
    stream = torch.cuda.current_stream().cuda_stream
    rc = lib.sglk_fp8_blockwise_scaled_grouped_mm_static(*_common_args(p, out), stream)
    assert rc == 0, lib.sglk_error_string(rc)
    return out


def _same(x, y):
    """Rows of experts: equal bits (and therefore inside the suite's tolerance); rows of no expert: untouched NaN in both."""
    torch.cuda.synchronize()
    written = ~torch.isnan(y.float())
    assert written.any()
    assert torch.equal(torch.isnan(x.float()), ~written), "the two entry points wrote different rows"
    xv, yv = torch.where(written, x, torch.zeros_like(x)), torch.where(written, y, torch.zeros_like(y))
    torch.testing.assert_close(xv, yv, rtol=1e-2, atol=5e-4)
    assert torch.equal(xv, yv)


@pytest.mark.gpu
def test_static_scratch_matches_workspace_form_on_consecutive_calls():
    lib = _lib.load()
    torch.cuda.set_device(0)
    # many single-CTA tiles (ragged, empty and one-row experts, 8 unused rows behind each) ...
    many = _problem([300, 0, 17, 129, 256, 1, 640, 90], 384, 1024, seed=1, gap=8)
    # ... then few 256-row pair tiles: the array still holds the tail of the longer table
    few = _problem([700, 513, 300], 512, 512, seed=2)
    want_many, want_few = _with_workspace(lib, many), _with_workspace(lib, few)
    _same(_static(lib, many), want_many)      # first call on this device in a fresh process: looks the address up
    _same(_static(lib, few), want_few)        # second call, right behind it, same array
    _same(_static(lib, many), want_many)
    # two calls queued on one stream without the host waiting in between (stream order is all they need)
    o1, o2 = _static(lib, few), _static(lib, many)
    _same(o1, want_few)
    _same(o2, want_many)


def test_static_scratch_rejects_problems_beyond_its_capacity():
    """Refused before any CUDA call, so this part also runs on a box without a GPU."""
    lib = _lib.load()
    cap = lib.sglk_grouped_mm_static_max_tiles()
    assert cap == (65536 - 16) // 16
    f = lib.sglk_fp8_blockwise_scaled_grouped_mm_static
    args = lambda E, rows: [16, 16, 16, 16, 16, 16, 16, E, rows, 256, 512, 512, 512, 256, 4, 1, 8, 2, 1, _lib.BF16, 0, None]
    assert f(*args(cap, 128)) == E_CAPACITY                       # cap experts + 1 tile
    assert f(*args(8, (cap - 7) * 128)) == E_CAPACITY             # cap - 7 tiles + 8 experts
    assert b"capacity" in lib.sglk_error_string(E_CAPACITY)
    assert f(*args(0, 0)) == 0                                    # empty problem: nothing to do, nothing looked up
