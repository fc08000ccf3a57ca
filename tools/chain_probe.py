#!/usr/bin/env python
"""Attribution probe for decode_chain_kernel (GPU box): times a Qwen3-shaped chain under the
QL_CHAIN_DEBUG skip flags (results are WRONG under any flag) and dumps the per-op globaltimer
trace of the first 16 ops.   python tools/chain_probe.py [0.6b|8b] [layers] [M]"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "nano-vllm-hw3_b200")]
import torch  # noqa: E402
from nanovllm import _ext  # noqa: E402

SH = {"0.6b": [(4096, 1024), (1024, 2048), (6144, 1024), (1024, 3072)],
      "8b": [(6144, 4096), (4096, 4096), (24576, 4096), (4096, 12288)]}
model = sys.argv[1] if len(sys.argv) > 1 else "0.6b"
layers = int(sys.argv[2]) if len(sys.argv) > 2 else 28
M = int(sys.argv[3]) if len(sys.argv) > 3 else 16
torch.manual_seed(0)
ops = []
x0 = torch.randn(M, SH[model][0][1], device="cuda").to(torch.bfloat16)
for li in range(layers):
    for j, (N, K) in enumerate(SH[model]):
        qw = torch.randint(-127, 128, (N, K), device="cuda", dtype=torch.int8)
        sw = torch.full((N,), 2e-4, device="cuda", dtype=torch.float32)
        ops.append((qw, sw, None, x0 if (li == 0 and j == 0) else None))
chain = _ext.DecodeChain(ops, M, torch.int8)
nbytes = sum(o[0].numel() for o in ops)


def timeit(n=20):
    for _ in range(3):
        chain.run()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n):
        chain.run()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n


ms = timeit()
print(f"{model} x{layers} layers M={M}: {ms * 1e3:9.1f} us/step  {ms * 1e3 / len(ops):7.2f} us/op  {nbytes / ms / 1e6:8.1f} GB/s", flush=True)

# clock64 trace of the first 16 ops (per CTA deltas, cycles)
G = torch.cuda.get_device_properties(0).multi_processor_count
lab = ["start", "amax sync", "quantised", "1st slot full", "pass-0 mma done", "partials landed", "pass-0 finished",
       "op finished", "pre-arrive sync", "arrived", "barrier done"]
tr = torch.zeros(G * 16 * 16, dtype=torch.int64, device="cuda")
_ext.load().ql_debug_set_trace(tr.data_ptr())
chain.run()
torch.cuda.synchronize()
_ext.load().ql_debug_set_trace(None)
t = tr.view(G, 16, 16).cpu().double()
print("per-op cycles from op start, median over stamping CTAs (max in brackets)")
print("stamps (mma.sync kernel): " + ", ".join(f"{i}={n}" for i, n in enumerate(lab)))
print("stamps (tcgen05 kernel): 0=start, 1=amax sync, 2=B operand quantised, 3=first tile accumulated, 4=first partials landed, "
      "7=op finished, 8=pre-arrive sync, 9=arrived, 10=barrier done, 11=MMA warp: B ready, 12/13=first / last K block of the "
      "first tile in shared memory, 14/15=TMA of those blocks issued (negative = before the op started)")
for l in range(min(16, len(ops))):
    row = []
    for i in range(1, 16):
        ok = (t[:, l, i] > 0) & (t[:, l, 0] > 0)
        v = (t[:, l, i] - t[:, l, 0])[ok]
        row.append("    -    " if v.numel() == 0 else f"{v.median():5.0f}[{v.max():5.0f}]")
    N, K = ops[l][0].shape
    print(f"op {l:2d} {N:5d}x{K:5d} " + " ".join(row))
