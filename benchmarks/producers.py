#!/usr/bin/env python
"""Fused producers vs the unfused sequence (bf16 producer + per-row quantiser), B200.

    python benchmarks/producers.py [--tokens 8192]

Algorithmic bytes per row: add+RMSNorm+quant = 2K (x) + 2K (residual in) + 2K (residual out) + K (q)
+ 4; SiluAndMul+quant = 4I (gate_up) + I (q) + 4.  Buffers are rotated over > 256 MB so the
figures are HBM figures (L2 is 126 MB).  CUDA events around a graph replay of `reps` launches.
"""
import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "nano-vllm-hw3_b200")):
    sys.path.insert(0, p)
import torch  # noqa: E402
from nanovllm import _ext  # noqa: E402


def timed(fn, nbuf, reps=3):
    s = torch.cuda.Stream()
    s.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(s):
        for i in range(nbuf):
            fn(i)
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g, stream=s):
            for i in range(nbuf):
                fn(i)
    torch.cuda.current_stream().wait_stream(s)
    for _ in range(3):
        g.replay()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        g.replay()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) * 1e3 / (reps * nbuf)      # us per call


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--tokens", type=int, default=8192)
    a = ap.parse_args()
    peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"] \
        if os.path.exists(os.path.join(ROOT, "MEASURED_PEAKS.json")) else 6650.0
    M = a.tokens
    print(f"# M={M} tokens, HBM peak {peaks} GB/s; us per call, GB/s = algorithmic bytes / time")
    print(f"{'op':28s} {'M':>6s} {'K/I':>6s} {'fused us':>9s} {'GB/s':>7s} {'frac':>6s} {'unfused us':>11s} {'speedup':>8s}")
    for dt in (torch.int8, torch.float8_e4m3fn):
        for K in (1024, 2048, 4096):
            nbuf = max(2, int(300e6 // (M * K * 7)) + 1)
            xs = [torch.randn(M, K, device="cuda").to(torch.bfloat16) for _ in range(nbuf)]
            rs = [torch.randn(M, K, device="cuda").to(torch.bfloat16) for _ in range(nbuf)]
            w = torch.ones(K, device="cuda", dtype=torch.bfloat16)
            t_f = timed(lambda i: _ext.rmsnorm_quant(xs[i], w, 1e-6, rs[i], dt), nbuf)

            def unfused(i):
                _, _, h = _ext.rmsnorm_quant(xs[i], w, 1e-6, rs[i], None)
                _ext.quant_rows(h, dt)
            t_u = timed(unfused, nbuf)
            by = M * (7 * K + 4)
            print(f"{'add+rmsnorm+quant ' + str(dt)[6:10]:28s} {M:6d} {K:6d} {t_f:9.2f} {by / t_f / 1e3:7.0f} "
                  f"{by / t_f / 1e3 / peaks:6.3f} {t_u:11.2f} {t_u / t_f:8.2f}")
            del xs, rs
        for I in (3072, 6144, 12288):
            nbuf = max(2, int(300e6 // (M * I * 5)) + 1)
            gus = [torch.randn(M, 2 * I, device="cuda").to(torch.bfloat16) for _ in range(nbuf)]
            t_f = timed(lambda i: _ext.silu_mul_quant(gus[i], dt), nbuf)

            def unfused(i):
                _, _, h = _ext.silu_mul_quant(gus[i], None)
                _ext.quant_rows(h, dt)
            t_u = timed(unfused, nbuf)
            by = M * (5 * I + 4)
            print(f"{'silu*mul+quant ' + str(dt)[6:10]:28s} {M:6d} {I:6d} {t_f:9.2f} {by / t_f / 1e3:7.0f} "
                  f"{by / t_f / 1e3 / peaks:6.3f} {t_u:11.2f} {t_u / t_f:8.2f}")
            del gus


if __name__ == "__main__":
    main()
