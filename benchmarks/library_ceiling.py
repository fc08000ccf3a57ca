#!/usr/bin/env python
"""GPU library ceiling: cuBLASLt INT8 (torch._int_mm) and FP8
(torch._scaled_mm) on 8192^3 and on the Qwen3 prefill shapes, next to this repo's GEMM on the same
shapes.  Library calls are the yardstick here, never the product path."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "nano-vllm-hw3_b200")):
    sys.path.insert(0, p)
import torch  # noqa: E402
from nanovllm import _ext  # noqa: E402


def t_us(fn, reps=20):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) * 1e3 / reps


print(f"{'shape (M,N,K)':26s} {'int_mm us':>10s} {'TF/s':>7s} {'scaled_mm us':>13s} {'TF/s':>7s} {'ql i8 us':>9s} {'TF/s':>7s} {'ql f8 us':>9s} {'TF/s':>7s}")
for M, N, K in [(8192, 8192, 8192), (8192, 24576, 4096), (8192, 4096, 12288), (4096, 4096, 1024), (4096, 6144, 1024),
                (4096, 1024, 2048), (4096, 1024, 3072)]:
    fl = 2.0 * M * N * K
    a8 = torch.randint(-127, 127, (M, K), device="cuda", dtype=torch.int8)
    b8 = torch.randint(-127, 127, (N, K), device="cuda", dtype=torch.int8)
    sx = torch.rand(M, device="cuda") + 0.5
    sw = torch.rand(N, device="cuda") + 0.5
    t_i = t_us(lambda: torch._int_mm(a8, b8.t()))
    af, bf = a8.float().to(torch.float8_e4m3fn), b8.float().to(torch.float8_e4m3fn)
    one = torch.ones((), device="cuda")
    try:
        t_f = t_us(lambda: torch._scaled_mm(af, bf.t(), scale_a=one, scale_b=one, out_dtype=torch.bfloat16))
    except Exception as e:  # noqa: BLE001
        t_f = float("nan")
    t_q = t_us(lambda: _ext.gemm_w8a8(a8, sx, b8, sw, None, torch.int8))
    t_q8 = t_us(lambda: _ext.gemm_w8a8(af, sx, bf, sw, None, torch.float8_e4m3fn))
    print(f"{str((M, N, K)):26s} {t_i:10.1f} {fl / t_i / 1e6:7.0f} {t_f:13.1f} {fl / t_f / 1e6:7.0f} "
          f"{t_q:9.1f} {fl / t_q / 1e6:7.0f} {t_q8:9.1f} {fl / t_q8 / 1e6:7.0f}")
