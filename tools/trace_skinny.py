#!/usr/bin/env python
"""Phase timeline of the decode kernel (gemm_skinny.cu) from its clock64() trace stamps.

    python tools/trace_skinny.py [--M 1] [--N 4096] [--K 1024] [--group 0] [--dtype int8]

Registers a 16 x int64 device buffer with ql_debug_set_trace, runs ql_linear a few times and
prints, for CTA 0, the cycle count of every phase relative to kernel entry (median of runs).
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


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--M", type=int, default=1)
    ap.add_argument("--N", type=int, default=4096)
    ap.add_argument("--K", type=int, default=1024)
    ap.add_argument("--group", type=int, default=0)
    ap.add_argument("--dtype", default="int8")
    ap.add_argument("--runs", type=int, default=9)
    a = ap.parse_args()
    from nanovllm import _ext
    dt = torch.int8 if a.dtype == "int8" else torch.float8_e4m3fn
    w = (torch.randn(a.N, a.K, device="cuda") * 0.02).to(torch.bfloat16)
    q, s = _ext.quant_groups(w, a.group, dt) if a.group else _ext.quant_rows(w, dt)
    x = torch.randn(a.M, a.K, device="cuda").to(torch.bfloat16)
    buf = torch.zeros(16, dtype=torch.int64, device="cuda")
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
    rows = []
    _ext.linear(x, q, s, None, dt, a.group)
    _ext.load().ql_debug_set_trace(buf.data_ptr())
    for _ in range(a.runs):
        flush.zero_()                      # weights come from HBM, as in a real decode step
        buf.zero_()
        torch.cuda.synchronize()
        _ext.linear(x, q, s, None, dt, a.group)
        torch.cuda.synchronize()
        rows.append(buf.cpu().tolist())
    _ext.load().ql_debug_set_trace(None)
    print(f"M={a.M} N={a.N} K={a.K} g={a.group} {a.dtype} path={_ext.linear_path(a.M, a.N, a.K, a.group)}")
    med = []
    for i in range(15):
        v = sorted(r[i] - r[0] for r in rows if r[i] != 0)
        med.append(v[len(v) // 2] if v else None)
    for i, name in enumerate(NAMES):
        if med[i] is not None:
            print(f"  {i:2d} {name:28s} {med[i]:8d} cyc  ({med[i] / 1.9:8.0f} ns @1.9GHz)")


if __name__ == "__main__":
    main()
