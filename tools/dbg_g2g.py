import os, sys
os.environ['QL_GROUPED_PAIR'] = '2'
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "nano-vllm-hw3_b200")]
import numpy as np, torch
from nanovllm import _ext
from oracle import oracle as orc
DT = {"i8": torch.int8, "f8": torch.float8_e4m3fn}; FMT = {"i8": orc.INT8, "f8": orc.FP8}
def rand_bf16(shape, seed, scale=1.0):
    return (torch.randn(*shape, generator=torch.Generator().manual_seed(seed)) * scale).to(torch.bfloat16)
for (M, N, K, g) in [(1000, 520, 2048, 256), (1000, 520, 2048, 128), (1024, 512, 2048, 256), (256, 256, 1024, 256), (256, 256, 2048, 512)]:
    for dn in ("i8", "f8"):
        x, w = rand_bf16((M, K), 1), rand_bf16((N, K), 2, 0.02)
        b = rand_bf16((N,), 31)
        qx, sx = orc.quant_rows(orc.t2n(x), FMT[dn]); qw, sw = orc.quant_groups(orc.t2n(w), g, FMT[dn])
        y = _ext.gemm_w8a8(orc.n2t(qx, DT[dn]).cuda(), torch.from_numpy(sx).cuda(), orc.n2t(qw, DT[dn]).cuda(),
                           torch.from_numpy(sw).cuda(), b.cuda(), DT[dn], g)
        yo = orc.gemm_groups(qx, sx, qw, sw, orc.t2n(b), g, FMT[dn])
        got = orc.bf16_decode(orc.t2n(y)); ref = orc.bf16_decode(yo)
        bad = np.argwhere(orc.t2n(y) != yo)
        print(M, N, K, g, dn, _ext.linear_path(M, N, K, g), "mismatches", len(bad), "max err", np.abs(got - ref).max(), "max ref", np.abs(ref).max())
        if len(bad):
            rows, cols = np.unique(bad[:, 0]), np.unique(bad[:, 1])
            print("   rows", rows[:10], "...", rows[-5:], len(rows), " cols", cols[:10], "...", cols[-5:], len(cols))
            i, j = bad[0]; print("   first", i, j, got[i, j], ref[i, j])
