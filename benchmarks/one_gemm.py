#!/usr/bin/env python
"""Launch one quantized-linear shape a few times (for ncu captures and quick timings).

    python benchmarks/one_gemm.py M N K [--dtype int8|fp8] [--group 0] [--reps 5] [--linear]
"""
import argparse
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for _p in (ROOT, os.path.join(ROOT, "nano-vllm-hw3_b200")):
    if _p not in sys.path:
        sys.path.insert(0, _p)
import torch  # noqa: E402
from nanovllm import _ext  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("M", type=int)
ap.add_argument("N", type=int)
ap.add_argument("K", type=int)
ap.add_argument("--dtype", default="int8")
ap.add_argument("--group", type=int, default=0)
ap.add_argument("--reps", type=int, default=5)
ap.add_argument("--linear", action="store_true", help="whole ql_linear (quantiser + GEMM / fused GEMV)")
a = ap.parse_args()
dt = torch.int8 if a.dtype == "int8" else torch.float8_e4m3fn
g = torch.Generator(device="cuda").manual_seed(0)
w = (torch.randn(a.N, a.K, device="cuda", generator=g) * 0.02).to(torch.bfloat16)
x = torch.randn(a.M, a.K, device="cuda", generator=g).to(torch.bfloat16)
qw, sw = _ext.quant_groups(w, a.group, dt) if a.group else _ext.quant_rows(w, dt)
qx, sx = _ext.quant_rows(x, dt)


def run():
    if a.linear or a.M <= 16:
        return _ext.linear(x, qw, sw, None, dt, a.group)
    return _ext.gemm_w8a8(qx, sx, qw, sw, None, dt, a.group)


for _ in range(3):
    run()
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(a.reps):
    run()
e1.record()
torch.cuda.synchronize()
t = e0.elapsed_time(e1) / a.reps * 1e-3
print(f"M={a.M} N={a.N} K={a.K} {a.dtype} g={a.group} path={_ext.linear_path(a.M, a.N, a.K, a.group)} "
      f"{t * 1e6:.1f} us  {2.0 * a.M * a.N * a.K / t / 1e12:.0f} TFLOP/s")
