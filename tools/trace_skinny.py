#!/usr/bin/env python
"""Phase timeline of the decode kernel (gemm_skinny.cu) from its per-CTA clock64() trace stamps.

    python tools/trace_skinny.py [--M 1] [--N 4096] [--K 1024] [--group 0] [--dtype int8] [--prequant]

Registers a [4096 x 16] int64 device buffer with ql_debug_set_trace, runs one ql_linear (or, with
--prequant, one ql_gemm_w8a8 on pre-quantised activations) after evicting L2 by reading a large
buffer, and prints for every phase the min / median / max over CTAs of the time since the
earliest CTA entered the kernel (globaltimer aligns CTAs across SMs, clock64 measures inside).
"""
import argparse
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for _p in (ROOT, os.path.join(ROOT, "nano-vllm-hw3_b200")):
    if _p not in sys.path:
        sys.path.insert(0, _p)
import torch  # noqa: E402

NAMES = ["entry", "setup done", "producer: ring primed", "quant: pdl_wait done", "quant: pass-1 amax",
         "quant: cluster amax", "quant: codes in smem", "mma: B ready", "mma: first stage full",
         "mma: all issued", "epi: accumulator ready", "epi: scattered", "epi: reduce barrier",
         "epi: stored", "exit"]
MAXG = 4096


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--M", type=int, default=1)
    ap.add_argument("--N", type=int, default=4096)
    ap.add_argument("--K", type=int, default=1024)
    ap.add_argument("--group", type=int, default=0)
    ap.add_argument("--dtype", default="int8")
    ap.add_argument("--prequant", action="store_true")
    ap.add_argument("--ghz", type=float, default=1.9)
    a = ap.parse_args()
    from nanovllm import _ext
    dt = torch.int8 if a.dtype == "int8" else torch.float8_e4m3fn
    w = (torch.randn(a.N, a.K, device="cuda") * 0.02).to(torch.bfloat16)
    q, s = _ext.quant_groups(w, a.group, dt) if a.group else _ext.quant_rows(w, dt)
    x = torch.randn(a.M, a.K, device="cuda").to(torch.bfloat16)
    qx, sx = _ext.quant_rows(x, dt)
    buf = torch.zeros(MAXG * 16, dtype=torch.int64, device="cuda")
    flush = torch.ones(512 << 20, dtype=torch.uint8, device="cuda")

    def run():
        if a.prequant:
            return _ext.gemm_w8a8(qx, sx, q, s, None, dt, a.group)
        return _ext.linear(x, q, s, None, dt, a.group)

    run()
    _ext.load().ql_debug_set_trace(buf.data_ptr())
    for _ in range(3):
        flush.sum()                        # evict by READING: no dirty lines compete with the kernel
        buf.zero_()
        x.add_(0)                          # bring the activations back into L2, as in a real decode step
        qx.add_(0)
        torch.cuda.synchronize()
        run()
        torch.cuda.synchronize()
    _ext.load().ql_debug_set_trace(None)
    t = buf.cpu().view(MAXG, 16)
    live = t[:, 0] != 0
    t = t[live]
    g0 = t[:, 15].min()
    print(f"M={a.M} N={a.N} K={a.K} g={a.group} {a.dtype} prequant={a.prequant} "
          f"path={_ext.linear_path(a.M, a.N, a.K, a.group)} CTAs={t.shape[0]}")
    entry_ns = (t[:, 15] - g0).double()
    print(f"  {'phase':30s} {'min':>8s} {'median':>8s} {'max':>8s}   (ns since the first CTA's entry)")
    for i, name in enumerate(NAMES):
        ok = t[:, i] != 0
        if not ok.any():
            continue
        v = entry_ns[ok] + (t[ok, i] - t[ok, 0]).double() / a.ghz
        print(f"  {i:2d} {name:27s} {v.min().item():8.0f} {v.median().item():8.0f} {v.max().item():8.0f}")


if __name__ == "__main__":
    main()
