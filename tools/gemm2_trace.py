#!/usr/bin/env python
"""Timeline of back-to-back gemm2_kernel launches (GPU box): 8 launches of one shape on distinct
weights, eager with PDL, globaltimer stamps per CTA.   python tools/gemm2_trace.py M N K"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "nano-vllm-hw3_b200")]
import torch
from nanovllm import _ext
M, N, K = (int(v) for v in sys.argv[1:4])
g = torch.Generator(device="cuda").manual_seed(0)
ws = [torch.randint(-127, 128, (N, K), device="cuda", dtype=torch.int8) for _ in range(8)]
sw = torch.full((N,), 1e-3, device="cuda")
qx = torch.randint(-127, 128, (M, K), device="cuda", dtype=torch.int8)
sx = torch.full((M,), 1e-2, device="cuda")
ys = [torch.empty(M, N, device="cuda", dtype=torch.bfloat16) for _ in range(8)]
def run():
    for w, y in zip(ws, ys):
        _ext.gemm_w8a8(qx, sx, w, sw, None, torch.int8, 0, out=y)
s = torch.cuda.Stream()
with torch.cuda.stream(s):
    run(); run()
    gr = torch.cuda.CUDAGraph()
    with torch.cuda.graph(gr, stream=s):
        run()
    torch.cuda.synchronize()
    gr.replay(); torch.cuda.synchronize()
    tr = torch.zeros(8 * 148 * 8, dtype=torch.int64, device="cuda")
    _ext.load().ql_debug_set_trace(tr.data_ptr())
    # the trace pointer is baked at launch time: re-capture with tracing on
    gr2 = torch.cuda.CUDAGraph()
    with torch.cuda.graph(gr2, stream=s):
        run()
    gr2.replay(); torch.cuda.synchronize()
    gr2.replay(); torch.cuda.synchronize()
    _ext.load().ql_debug_set_trace(None)
t = tr.view(8, 148, 8).cpu().double()
lab = ["entry", "setup done", "producer past pdl wait", "first MMA", "last MMA issued", "last acc ready", "epilogue done", "exit"]
order = sorted(range(8), key=lambda i: t[i][t[i][:, 0] > 0][:, 0].min().item() if (t[i][:, 0] > 0).any() else 0)
base = min(t[i][t[i][:, 0] > 0][:, 0].min().item() for i in order if (t[i][:, 0] > 0).any())
print(f"shape M={M} N={N} K={K}; ns since the first CTA entry of the first launch; min/median/max over CTAs")
print("stamps:", ", ".join(f"{i}={n}" for i, n in enumerate(lab)))
for i in order:
    row = []
    for j in range(8):
        v = t[i][:, j]; v = v[v > 0] - base
        row.append("   -   " if v.numel() == 0 else f"{v.min():6.0f}/{v.median():6.0f}/{v.max():6.0f}")
    print(f"launch slot {i}: " + " | ".join(row))
