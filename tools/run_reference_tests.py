#!/usr/bin/env python
"""Run the REFERENCE's own test files, unmodified, against this package (drop-in check).

The reference's tests are device-agnostic (tests/utils.py picks "cuda" when it is there) and import `sgl_kernel` — the
name of this repo's package — so with sgl-kernel-xpu_b200/ first on PYTHONPATH they exercise this library through the
reference's own call sites:
    tests/test_fp8_blockwise_gemm.py            576 cases (M x N x K x out_dtype grid, :88-91)
    tests/test_fp8_blockwise_moe.py             grouped GEMM, E in {8, 16} x {half, bfloat16}
    tests/test_per_token_group_quant_8bit.py    quant v1 grid
    tests/test_per_token_quant_fp8.py
Needs a checkout of the reference (SGLK_REFERENCE=/path/to/sgl-kernel-xpu) AND a GPU on the same machine; nothing is
copied from it.  The GPU boxes of this project have no checkout, so there the same grids run from tests/ instead
(test_gemm_gpu.py::test_reference_grid_*, test_quant_gpu.py::test_v2_reference_grid, test_grouped_gpu.py), with the
reference's own baselines restated in oracle/.  Log: gpurun_out/reference_tests.log."""
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
FILES = ["test_fp8_blockwise_gemm.py", "test_fp8_blockwise_moe.py", "test_per_token_group_quant_8bit.py",
         "test_per_token_quant_fp8.py"]


def main():
    ref = os.environ.get("SGLK_REFERENCE", "")
    tests = os.path.join(ref, "tests")
    if not ref or not os.path.isdir(tests):
        print("SGLK_REFERENCE does not point at a reference checkout: nothing to run")
        return 2
    import torch

    if not torch.cuda.is_available():
        print("no CUDA device: the reference's tests would run its CPU fallbacks against nothing — not run")
        return 2
    env = dict(os.environ)
    env["PYTHONPATH"] = os.pathsep.join([os.path.join(ROOT, "sgl-kernel-xpu_b200"), tests, env.get("PYTHONPATH", "")])
    os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
    log = os.path.join(ROOT, "gpurun_out", "reference_tests.log")
    rc = 0
    with open(log, "w") as f:
        for name in FILES:
            path = os.path.join(tests, name)
            if not os.path.exists(path):
                f.write(f"== {name}: not in this checkout\n")
                continue
            f.write(f"== {name}\n")
            f.flush()
            r = subprocess.run([sys.executable, "-m", "pytest", path, "-q", "-x", "-p", "no:cacheprovider"], env=env, cwd=tests,
                               stdout=f, stderr=subprocess.STDOUT)
            rc = rc or r.returncode
    print(open(log).read()[-2000:])
    return rc


if __name__ == "__main__":
    sys.exit(main())
