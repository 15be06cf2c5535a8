"""CPU oracle for the FP8 block-quantized linear path — TEST INFRASTRUCTURE ONLY.

Importable from `tests/`, `__graft_entry__.smoke()` and `bench.py`'s CPU-baseline legs; the
product package (`sgl-kernel-xpu_b200/`) must never import it.  See torch_oracle.py / quant_oracle.c.
"""
