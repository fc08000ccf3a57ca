#!/usr/bin/env python
"""Quantized LM head (vocab 151 936 x hidden) decode timing vs the HBM roofline, B200.
Weights: N*K bytes (int8/fp8) stream once per call; 3 distinct copies rotate (> 126 MB L2)."""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "nano-vllm-hw3_b200")):
    sys.path.insert(0, p)
import torch  # noqa: E402
from nanovllm import _ext  # noqa: E402
from benchmarks.producers import timed  # noqa: E402

pk = os.path.join(ROOT, "MEASURED_PEAKS.json")
peak = json.load(open(pk))["hbm_gbs"] if os.path.exists(pk) else 6650.0
V = 151936
print(f"# lm_head [V={V}, H] per-row, whole ql_linear (quantiser + decode kernel), HBM peak {peak} GB/s")
print(f"{'H':>6s} {'dtype':>6s} {'M':>4s} {'us':>8s} {'GB/s':>7s} {'frac':>6s}  path")
for H in (1024, 4096):
    for dt in (torch.int8, torch.float8_e4m3fn):
        ws = []
        for i in range(3):
            w = torch.randn(V, H, device="cuda", dtype=torch.bfloat16) * 0.02
            ws.append(_ext.quant_rows(w, dt))
            del w
        for M in (1, 16, 64, 256):
            x = torch.randn(M, H, device="cuda").to(torch.bfloat16)
            t = timed(lambda i: _ext.linear(x, ws[i][0], ws[i][1], None, dt, 0), 3, reps=5)
            by = V * H + 4 * V + 2 * M * H + 2 * M * V
            print(f"{H:6d} {str(dt)[6:10]:>6s} {M:4d} {t:8.2f} {by / t / 1e3:7.0f} {by / t / 1e3 / peak:6.3f}  "
                  f"{_ext.linear_path(M, V, H)}")
        del ws
