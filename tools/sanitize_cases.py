#!/usr/bin/env python
"""A reduced selection of the GPU parity cases for compute-sanitizer:

    compute-sanitizer --tool memcheck|racecheck|synccheck python tools/sanitize_cases.py

One small problem per kernel family -- gemm2_kernel (CTA pair, int8 + fp8, bias, M/N tails),
gemm_kernel (per-row + per-group), skinny_kernel (S = 1, 2, 8; pre-quantised and fused quantiser;
per-group), the quantisers and the fused producers -- each checked against the CPU oracle, so a
sanitizer run also proves the instrumented kernels still compute the right thing.  Sizes are
tiny on purpose: the sanitizer slows kernels down by two to three orders of magnitude.
Exit code 0 only if every case matched.
"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "nano-vllm-hw3_b200")):
    sys.path.insert(0, p)

import numpy as np  # noqa: E402
import torch  # noqa: E402

from nanovllm import _ext  # noqa: E402
from oracle import oracle as orc  # noqa: E402

DT = {orc.INT8: torch.int8, orc.FP8: torch.float8_e4m3fn}


def rnd(shape, seed, scale=1.0):
    return (torch.randn(*shape, generator=torch.Generator().manual_seed(seed)) * scale).to(torch.bfloat16)


def close(y, yo, exact):
    if exact:
        return np.array_equal(y, yo)
    a, b = orc.bf16_decode(y).astype(np.float64), orc.bf16_decode(yo).astype(np.float64)
    return np.abs(a - b).max() <= 2.0 ** (np.floor(np.log2(np.abs(b).max())) - 7)


def linear_case(name, M, N, K, fmt, group=0, bias=False, env=None, gemm_only=False):
    for k, v in (env or {}).items():
        os.environ[k] = v
    x, w = rnd((M, K), M + N), rnd((N, K), N + K, 0.02)
    b = rnd((N,), 5) if bias else None
    qw, sw = orc.quant_groups(orc.t2n(w), group, fmt) if group else orc.quant_rows(orc.t2n(w), fmt)
    qx, sx = orc.quant_rows(orc.t2n(x), fmt)
    bo = None if b is None else orc.t2n(b)
    yo = orc.gemm_groups(qx, sx, qw, sw, bo, group, fmt) if group else orc.gemm_rows(qx, sx, qw, sw, bo, fmt)
    dqw, dsw = orc.n2t(qw, DT[fmt]).cuda(), torch.from_numpy(sw).cuda()
    bc = None if b is None else b.cuda()
    if gemm_only:
        y = _ext.gemm_w8a8(orc.n2t(qx, DT[fmt]).cuda(), torch.from_numpy(sx).cuda(), dqw, dsw, bc, DT[fmt], group)
    else:
        y = _ext.linear(x.cuda(), dqw, dsw, bc, DT[fmt], group)
    torch.cuda.synchronize()
    path = _ext.linear_path(M, N, K, group)
    ok = close(orc.t2n(y), yo, fmt == orc.INT8 and not (group and path.startswith(("skinny", "quant+skinny"))))
    for k in (env or {}):
        os.environ.pop(k, None)
    print(f"{'ok  ' if ok else 'FAIL'} {name}: M={M} N={N} K={K} fmt={fmt} g={group} path={path}", flush=True)
    return ok


def main():
    assert torch.cuda.is_available()
    ok = True
    # CTA-pair GEMM (gemm2_kernel)
    ok &= linear_case("gemm2 int8 tails+bias", 300, 264, 256, orc.INT8, bias=True)
    ok &= linear_case("gemm2 fp8", 256, 512, 128, orc.FP8, gemm_only=True)
    # single-CTA GEMM (gemm_kernel): per-row short-K mid-M, per-group prefill
    ok &= linear_case("gemm per-row", 64, 256, 256, orc.INT8)
    ok &= linear_case("gemm per-group int8", 200, 136, 256, orc.INT8, group=128, bias=True)
    ok &= linear_case("gemm per-group fp8", 130, 128, 256, orc.FP8, group=128)
    # decode kernel (skinny_kernel): split factors, fused quantiser, per-group
    for S in ("1", "2", "8"):
        ok &= linear_case(f"skinny S={S}", 7, 264, 1024, orc.INT8, bias=True,
                          env={"QL_SKINNY_S": S, "QL_SKINNY_ALWAYS": "1"})
        ok &= linear_case(f"skinny fused S={S}", 5, 136, 1024, orc.FP8,
                          env={"QL_SKINNY_S": S, "QL_SKINNY_ALWAYS": "1", "QL_FUSED": "1"})
    ok &= linear_case("skinny M=40 S=auto", 40, 256, 1152, orc.INT8)
    ok &= linear_case("skinny per-group", 16, 256, 512, orc.FP8, group=128)
    # raw accumulators through both GEMM kernels
    g = torch.Generator().manual_seed(1)
    for M, N, K in ((130, 136, 256), (9, 40, 384)):
        qx = torch.randint(-127, 128, (M, K), generator=g, dtype=torch.int8)
        qw = torch.randint(-127, 128, (N, K), generator=g, dtype=torch.int8)
        acc = _ext.gemm_i8_acc(qx.cuda(), qw.cuda()).cpu().numpy()
        good = np.array_equal(acc, orc.gemm_i8_acc(qx.numpy(), qw.numpy()))
        print(f"{'ok  ' if good else 'FAIL'} i8 acc M={M}", flush=True)
        ok &= good
    # quantisers
    for fmt in (orc.INT8, orc.FP8):
        x = rnd((9, 3072), 3)
        q, s = _ext.quant_rows(x.cuda(), DT[fmt])
        qo, so = orc.quant_rows(orc.t2n(x), fmt)
        good = np.array_equal(orc.t2n(q), qo) and np.array_equal(orc.t2n(s), so)
        w = rnd((24, 512), 4, 0.02)
        q, s = _ext.quant_groups(w.cuda(), 128, DT[fmt])
        qo, so = orc.quant_groups(orc.t2n(w), 128, fmt)
        good &= np.array_equal(orc.t2n(q), qo) and np.array_equal(orc.t2n(s), so)
        q, sc = _ext.quant_tensor(w.cuda(), DT[fmt])
        qo, sco = orc.quant_tensor(orc.t2n(w), fmt)
        good &= np.array_equal(orc.t2n(q), qo) and float(sc) == sco
        print(f"{'ok  ' if good else 'FAIL'} quantisers fmt={fmt}", flush=True)
        ok &= good
    # persistent decode chain (clusters, DSMEM reduce-scatter, TMA ring, grid barrier): 4 chained ops,
    # K splits 4 and 2, an op with its own input, bias
    g2 = torch.Generator().manual_seed(2)
    shapes, M = [(264, 384), (128, 256), (1032, 128), (8, 1024)], 5
    ops, xs = [], []
    for i, (N, K) in enumerate(shapes):
        w = (torch.randn(N, K, generator=g2) * 0.05).to(torch.bfloat16)
        qw, sw = _ext.quant_rows(w.cuda(), torch.int8)
        b = (torch.randn(N, generator=g2) * 0.1).to(torch.bfloat16).cuda() if i in (0, 2) else None
        x = torch.randn(M, K, generator=g2).to(torch.bfloat16).cuda() if i in (0, 2) else None
        ops.append((qw, sw, b, x))
        xs.append(x)
    chain = _ext.DecodeChain(ops, M, torch.int8)
    chain.run()
    chain.run()
    torch.cuda.synchronize()
    good = True
    for i, (qw, sw, b, _) in enumerate(ops):
        xin = xs[i] if xs[i] is not None else chain.outputs[i - 1][:, :qw.shape[1]].contiguous()
        qx, sx = orc.quant_rows(orc.t2n(xin), orc.INT8)
        yo = orc.gemm_rows(qx, sx, orc.t2n(qw), orc.t2n(sw), None if b is None else orc.t2n(b), orc.INT8)
        good &= np.array_equal(orc.t2n(chain.outputs[i]), yo)
    print(f"{'ok  ' if good else 'FAIL'} decode chain", flush=True)
    ok &= good
    # fused producers
    x, r, wn = rnd((6, 1024), 7), rnd((6, 1024), 8), (1 + 0.1 * torch.randn(1024)).to(torch.bfloat16)
    res = r.cuda()
    q, s, h = _ext.rmsnorm_quant(x.cuda(), wn.cuda(), 1e-6, res, torch.int8, want_bf16=True)
    qo, so = orc.quant_rows(orc.t2n(h), orc.INT8)
    _, reso = orc.ref_rmsnorm(x, wn, 1e-6, r)
    good = np.array_equal(orc.t2n(q), qo) and np.array_equal(orc.t2n(s), so) and torch.equal(res.cpu(), reso)
    gu = rnd((6, 2 * 3072), 9)
    q, s, h = _ext.silu_mul_quant(gu.cuda(), torch.float8_e4m3fn, want_bf16=True)
    qo, so = orc.quant_rows(orc.t2n(h), orc.FP8)
    good &= np.array_equal(orc.t2n(q), qo) and np.array_equal(orc.t2n(s), so)
    print(f"{'ok  ' if good else 'FAIL'} producers", flush=True)
    ok &= good
    torch.cuda.synchronize()
    print("ALL OK" if ok else "FAILURES", flush=True)
    sys.exit(0 if ok else 1)


if __name__ == "__main__":
    main()
