"""Generate tests/golden/ref_golden.npz by RUNNING THE REFERENCE in the build container.

Run (here only -- /root/reference does not exist on the GPU box):

    cd /root/repo && python tests/golden/make_golden.py

What is executed (all reference code, unmodified, imported from /root/reference):

* torch functions on CPU tensors: ``per_group_quant`` (quantization.py:194-234),
  ``per_group_matmul`` (:265-307), ``per_tensor_quant`` (:166-191), ``per_tensor_matmul``
  (:237-262).
* the Triton kernels ``per_row_quant`` (:147-163) and ``per_row_matmul`` (:310-346) under
  ``TRITON_INTERPRET=1`` (numpy execution of the kernel source).

Harness shims: a stub ``torchao`` module (quantization.py:10 imports
it; it is not installed), and -- for the INT8 Triton path only -- ``libdevice.round`` is
replaced, in interpreter mode, by its definition (round half away from zero) written with
``tl`` ops, because the interpreter cannot call the CUDA libdevice extern.  Everything else
is the reference's own code path.

The file stores bit patterns: bf16 as uint16, fp8 as uint8.  Inputs are stored alongside
outputs so the tests do not depend on torch's RNG stream.
"""
import os
import sys
import types

os.environ["TRITON_INTERPRET"] = "1"
os.environ.setdefault("CUDA_VISIBLE_DEVICES", "")

REF = "/root/reference"
sys.path.insert(0, REF)

tao = types.ModuleType("torchao")
taoq = types.ModuleType("torchao.quantization")
for _n in ("quantize_", "Int8DynamicActivationInt8WeightConfig",
           "Float8DynamicActivationFloat8WeightConfig"):
    setattr(taoq, _n, None)
tao.quantization = taoq
sys.modules["torchao"] = tao
sys.modules["torchao.quantization"] = taoq

import numpy as np  # noqa: E402
import torch  # noqa: E402
import triton.language as tl  # noqa: E402
from triton.language.extra import libdevice  # noqa: E402


def _round_half_away(x):
    t = x.to(tl.int32).to(tl.float32)          # truncation toward zero (exact for |x| <= 127.5)
    d = tl.abs(x - t)                          # exact
    step = tl.where(x >= 0, 1.0, -1.0)
    return tl.where(d >= 0.5, t + step, t)


libdevice.round = _round_half_away

import nanovllm.utils.quantization as rq  # noqa: E402  (the reference)

assert rq.__file__.startswith(REF), rq.__file__

# the interpreter refuses CPU tensors unless told the "device" is fine; it accepts them as is.


def bits(t: torch.Tensor) -> np.ndarray:
    t = t.detach().cpu().contiguous()
    if t.dtype == torch.bfloat16:
        return t.view(torch.int16).numpy().view(np.uint16).copy()
    if t.dtype == torch.float8_e4m3fn:
        return t.view(torch.uint8).numpy().copy()
    if t.dtype == torch.int8:
        return t.numpy().view(np.uint8).copy()
    return t.numpy().copy()


def main():
    out = {}
    g = torch.Generator().manual_seed(20261104)

    def randn(*shape, scale=1.0):
        return (torch.randn(*shape, generator=g) * scale).to(torch.bfloat16)

    dts = {"i8": torch.int8, "f8": torch.float8_e4m3fn}

    # ---- A/B: per_group_quant + per_group_matmul (torch, CPU) ---------------------------
    for tag, (N, K, M) in {"a": (24, 256, 5), "b": (16, 512, 1)}.items():
        w = randn(N, K, scale=0.02)
        w[3].zero_()                       # exercises clamp(min=1e-8)
        w[5, :64] *= 37.0                  # one loud group
        x = randn(M, K)
        b = randn(N)
        out[f"pg_{tag}_w"] = bits(w)
        out[f"pg_{tag}_x"] = bits(x)
        out[f"pg_{tag}_bias"] = bits(b)
        for dn, dt in dts.items():
            for gs in (64, 128, K):
                qw, sw = rq.per_group_quant(w, gs, dt)
                y = rq.per_group_matmul(x, qw, sw, b, gs, dt)
                y_nb = rq.per_group_matmul(x.view(1, M, K), qw, sw, None, gs, dt)
                assert y_nb.shape == (1, M, N)
                out[f"pg_{tag}_{dn}_g{gs}_q"] = bits(qw)
                out[f"pg_{tag}_{dn}_g{gs}_s"] = bits(sw)
                out[f"pg_{tag}_{dn}_g{gs}_y"] = bits(y)
                out[f"pg_{tag}_{dn}_g{gs}_ynb"] = bits(y_nb.reshape(M, N))

    # ---- C: per_tensor_quant + per_tensor_matmul (torch, CPU) ---------------------------
    w = randn(32, 128, scale=0.05)
    x = randn(4, 128)
    b = randn(32)
    out["pt_w"], out["pt_x"], out["pt_bias"] = bits(w), bits(x), bits(b)
    for dn, dt in dts.items():
        qw, sw = rq.per_tensor_quant(w, dt)
        assert sw.dtype == torch.bfloat16 and sw.dim() == 0
        y = rq.per_tensor_matmul(x, qw, sw, b, dt)
        out[f"pt_{dn}_q"] = bits(qw)
        out[f"pt_{dn}_s"] = np.array([float(sw)], np.float32)
        out[f"pt_{dn}_y"] = bits(y)

    # ---- D/E: Triton per_row_quant / per_row_matmul under the interpreter ---------------
    for tag, (M, N, K) in {"a": (5, 136, 256), "b": (3, 8, 384), "c": (130, 24, 128)}.items():
        x = randn(M, K)
        if M > 2:
            x[1].zero_()                   # amax clamp 1e-8
            x[2] *= 1e-3
        w = randn(N, K, scale=0.02)
        out[f"pr_{tag}_x"], out[f"pr_{tag}_w"] = bits(x), bits(w)
        for dn, dt in dts.items():
            qx, sx = rq.per_row_quant(x, dt)
            qw, sw = rq.per_row_quant(w, dt)
            y = rq.per_row_matmul(x, qw, sw, None, dt)
            out[f"pr_{tag}_{dn}_qx"], out[f"pr_{tag}_{dn}_sx"] = bits(qx), bits(sx)
            out[f"pr_{tag}_{dn}_qw"], out[f"pr_{tag}_{dn}_sw"] = bits(qw), bits(sw)
            out[f"pr_{tag}_{dn}_y"] = bits(y)

    # ---- F: long-K INT8 case: accumulators beyond 2^24, where tl.cast(acc, fp32) (:111) rounds ----
    # K = 2048 with same-sign, near-full-scale rows: |acc| ~ 2048 * 120^2 = 2.9e7 > 2^24 and odd,
    # so the int32 -> fp32 conversion is inexact and its rounding mode is pinned by this fixture.
    M, N, K = 6, 24, 2048
    x = randn(M, K)
    w = randn(N, K, scale=0.02)
    x[0] = 3.0                                                   # q = 127 everywhere
    x[1] = (0.9 + 0.1 * torch.rand(K, generator=g)).to(torch.bfloat16) * 3.0
    x[2] = -(0.9 + 0.1 * torch.rand(K, generator=g)).to(torch.bfloat16) * 2.0
    w[0] = 0.02
    w[1] = (0.9 + 0.1 * torch.rand(K, generator=g)).to(torch.bfloat16) * 0.02
    w[2] = -(0.9 + 0.1 * torch.rand(K, generator=g)).to(torch.bfloat16) * 0.03
    out["pr_d_x"], out["pr_d_w"] = bits(x), bits(w)
    qx, sx = rq.per_row_quant(x, torch.int8)
    qw, sw = rq.per_row_quant(w, torch.int8)
    y = rq.per_row_matmul(x, qw, sw, None, torch.int8)
    acc = qx[:3].int() @ qw[:3].int().T
    assert int(acc.abs().max()) > (1 << 24) and bool((acc.abs() > (1 << 24)).sum() >= 8)
    assert bool(((acc.abs() > (1 << 24)) & (acc % 2 != 0)).any()), "need an accumulator fp32 cannot hold"
    out["pr_d_i8_qx"], out["pr_d_i8_sx"] = bits(qx), bits(sx)
    out["pr_d_i8_qw"], out["pr_d_i8_sw"] = bits(qw), bits(sw)
    out["pr_d_i8_y"] = bits(y)

    dst = os.path.join(os.path.dirname(os.path.abspath(__file__)), "ref_golden.npz")
    np.savez_compressed(dst, **out)
    print(f"wrote {dst}: {len(out)} arrays, {os.path.getsize(dst)} bytes")


if __name__ == "__main__":
    main()
