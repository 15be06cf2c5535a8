#!/usr/bin/env python
"""Run the REFERENCE's own test files, unmodified, against this package (drop-in check).

The reference's tests are device-agnostic (tests/utils.py picks "cuda" when it is there) and import `sgl_kernel` — the
name of this repo's package — so with sgl-kernel-xpu_b200/ first on PYTHONPATH they exercise this library through the
reference's own call sites:
    tests/test_fp8_blockwise_gemm.py            576 cases (M x N x K x out_dtype grid, :88-91)
    tests/test_fp8_blockwise_moe.py             grouped GEMM, E in {8, 16} x {half, bfloat16}
    tests/test_per_token_group_quant_8bit.py    quant v1 grid
    tests/test_per_token_quant_fp8.py
Needs the reference's tests/ directory (SGLK_REFERENCE=/path/whose/tests/ holds them) AND a GPU on the same machine.  The
GPU boxes of this project have no checkout: for a run there the five files + conftest.py + utils.py are copied into the
untracked, git-ignored scratch directory _ref_scratch/tests/ just before `gpurun` and deleted afterwards (results:
profiles/reference_tests_r03.txt).  Day to day the same grids run from tests/ instead
(test_gemm_gpu.py::test_reference_grid_*, test_quant_gpu.py::test_v2_reference_grid, test_grouped_gpu.py), with the
reference's own baselines restated in oracle/.  Log: gpurun_out/reference_tests.log."""
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
FILES = ["test_fp8_blockwise_gemm.py", "test_fp8_blockwise_moe.py", "test_per_token_group_quant_8bit.py",
         "test_per_token_group_quant_8bit_v2.py",
         "test_per_token_quant_fp8.py"]


RETARGET = [('"xpu"', '"cuda"'), ("torch.xpu.", "torch.cuda."), ("\n    device = device\n", "\n")]


def run_files(tests, env, f, label):
    rc = 0
    for name in FILES:
        path = os.path.join(tests, name)
        if not os.path.exists(path):
            f.write(f"== [{label}] {name}: not in this checkout\n")
            continue
        f.write(f"== [{label}] {name}\n")
        f.flush()
        r = subprocess.run([sys.executable, "-m", "pytest", path, "-q", "-p", "no:cacheprovider", "--tb=line"], env=env,
                           cwd=tests, stdout=f, stderr=subprocess.STDOUT)
        rc = rc or r.returncode
    return rc


def main():
    """Pass 1: the files exactly as they are.  Pass 2 (some of them name the device "xpu" literally, and
    test_fp8_blockwise_moe.py:67 shadows its module-level `device` with `device = device`, which raises on any machine):
    copies in a temporary directory with ONLY the substitutions of RETARGET applied — the device's name, nothing about
    what is called or compared.  Both passes are logged with their pass counts."""
    ref = os.environ.get("SGLK_REFERENCE", "")
    tests = os.path.join(ref, "tests")
    if not ref or not os.path.isdir(tests):
        print("SGLK_REFERENCE does not point at a reference checkout: nothing to run")
        return 2
    import torch

    if not torch.cuda.is_available():
        print("no CUDA device: the reference's tests would run its CPU fallbacks against nothing — not run")
        return 2
    os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
    log = os.path.join(ROOT, "gpurun_out", "reference_tests.log")

    def env_for(d):
        env = dict(os.environ)
        env["PYTHONPATH"] = os.pathsep.join([os.path.join(ROOT, "sgl-kernel-xpu_b200"), d, env.get("PYTHONPATH", "")])
        return env

    import tempfile

    with open(log, "w") as f, tempfile.TemporaryDirectory() as tmp:
        rc1 = run_files(tests, env_for(tests), f, "unmodified")
        for name in os.listdir(tests):
            if name.endswith(".py"):
                src = open(os.path.join(tests, name)).read()
                for a, b in RETARGET:
                    src = src.replace(a, b)
                open(os.path.join(tmp, name), "w").write(src)
        f.write(f"\n#### second pass: device name retargeted, substitutions = {RETARGET!r}\n")
        f.flush()
        rc2 = run_files(tmp, env_for(tmp), f, "retargeted")
    text = open(log).read()
    print("\n".join(ln for ln in text.splitlines() if ln.startswith("== ") or " passed" in ln or " failed" in ln or ln.startswith("FAILED")))
    return rc2


if __name__ == "__main__":
    sys.exit(main())
