import sys, os, torch
sys.path.insert(0, "nano-vllm-hw3_b200"); sys.path.insert(0, ".")
from nanovllm import _ext
from benchmarks.producers import timed
I = int(sys.argv[1]); M = 8192
nbuf = max(2, int(300e6 // (M * I * 5)) + 1)
gus = [torch.randn(M, 2 * I, device="cuda").to(torch.bfloat16) for _ in range(nbuf)]
for dt in (torch.int8, torch.float8_e4m3fn):
    t = timed(lambda i: _ext.silu_mul_quant(gus[i], dt), nbuf)
    print(os.environ.get("QL_PROD_CFG"), I, dt, f"{t:.2f} us  {M*(5*I+4)/t/1e3:.0f} GB/s")
