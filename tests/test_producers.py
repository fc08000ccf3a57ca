"""Fused producers (RMSNorm -> quant, SiluAndMul -> quant).

CPU part: the oracle restatements (oracle.ref_rmsnorm / ref_silu_and_mul) against goldens made
by running the reference's own modules (tests/golden/make_golden_producers.py), and argument
validation of the two C-ABI entry points.

GPU part (bars):
  * (q, s) of the fused kernel == ql_quant_rows applied to the kernel's own bf16 output h:
    BIT-EXACT (the quantiser arithmetic is the pinned per-row quantiser, only its input is fused);
  * residual_out == bf16(float(x) + float(residual)): BIT-EXACT;
  * h vs the reference golden / oracle: floating point (reduction order, exp implementation):
    every element within PROD_TOL_ULP (SiLU) / RMS_TOL_ULP (RMSNorm) bf16 ulp, at most PROD_MISMATCH of the elements differ at all;
  * feeding the fused (q, s) to the GEMM == ql_linear on h: BIT-EXACT.
"""
import ctypes
import os

import numpy as np
import pytest
import torch

from oracle import oracle as orc

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PROD_TOL_ULP = 1
RMS_TOL_ULP = 2        # a 1-ulp flip of bf16(x * rsqrt(var)) is scaled by |weight| (up to ~1.8 here)
PROD_MISMATCH = 2e-3
DT = {"i8": torch.int8, "f8": torch.float8_e4m3fn}


@pytest.fixture(scope="module")
def pgold():
    return np.load(os.path.join(ROOT, "tests", "golden", "ref_producers.npz"))


def bf(a):
    return orc.n2t(a, torch.bfloat16)


def ulp_diff(a_u16, b_u16):
    def key(u):
        u = u.astype(np.int32)
        return np.where(u & 0x8000, 0x8000 - u, u)
    return np.abs(key(a_u16) - key(b_u16))


def assert_bf16_close(got_u16, ref_u16, what, tol=PROD_TOL_ULP):
    d = ulp_diff(got_u16, ref_u16)
    # values below 2^-100 in magnitude (silu of a < -87: ex2.approx flushes the subnormal exp) are
    # compared absolutely: their ulp distance to 0 is large but they are 0 for any bf16 consumer
    tiny = np.abs(orc.bf16_decode(got_u16) - orc.bf16_decode(ref_u16)) <= 2.0 ** -100
    d = np.where(tiny, 0, d)
    assert d.max() <= tol, f"{what}: {d.max()} bf16 ulp"
    frac = (d > 0).mean()
    assert frac <= PROD_MISMATCH, f"{what}: {frac:.2e} of the elements differ"


# ------------------------------------------------------------------ CPU: oracle vs reference goldens

@pytest.mark.parametrize("tag", ["a", "b", "c"])
def test_oracle_rmsnorm_matches_reference(pgold, tag):
    x, r, w = (bf(pgold[f"rms_{tag}_{k}"]) for k in ("x", "r", "w"))
    h, _ = orc.ref_rmsnorm(x, w, 1e-6)
    assert np.array_equal(orc.t2n(h), pgold[f"rms_{tag}_h"])
    h, res = orc.ref_rmsnorm(x, w, 1e-6, r)
    assert np.array_equal(orc.t2n(h), pgold[f"rms_{tag}_hres"])
    assert np.array_equal(orc.t2n(res), pgold[f"rms_{tag}_res"])


@pytest.mark.parametrize("tag", ["a", "b", "c"])
def test_oracle_silu_matches_reference(pgold, tag):
    h = orc.ref_silu_and_mul(bf(pgold[f"silu_{tag}_x"]))
    assert np.array_equal(orc.t2n(h), pgold[f"silu_{tag}_h"])


def test_producer_argument_validation():
    from nanovllm import _ext
    lib = _ext.load()
    fake = ctypes.c_void_p(0x1000)
    f32 = ctypes.c_float(1e-6)
    # bad format
    assert lib.ql_rmsnorm_quant(fake, 128, None, None, fake, f32, 4, 128, 9, fake, fake, None, None) == 1
    assert lib.ql_silu_mul_quant(fake, 256, 4, 128, 9, fake, fake, None, None) == 1
    # K % 8 != 0, ld too small
    assert lib.ql_rmsnorm_quant(fake, 132, None, None, fake, f32, 4, 132, 0, fake, fake, None, None) == 2
    assert lib.ql_silu_mul_quant(fake, 128, 4, 128, 0, fake, fake, None, None) == 2
    # residual_in without residual_out, no output at all, q without s
    assert lib.ql_rmsnorm_quant(fake, 128, fake, None, fake, f32, 4, 128, 0, fake, fake, None, None) == 7
    assert lib.ql_rmsnorm_quant(fake, 128, None, None, fake, f32, 4, 128, 0, None, None, None, None) == 7
    assert lib.ql_silu_mul_quant(fake, 256, 4, 128, 0, fake, None, None, None) == 7
    # misaligned input
    assert lib.ql_silu_mul_quant(ctypes.c_void_p(0x1002), 256, 4, 128, 0, fake, fake, None, None) == 3
    # row too long for shared memory
    assert lib.ql_rmsnorm_quant(fake, 1 << 20, None, None, fake, f32, 4, 1 << 20, 0, fake, fake, None, None) == 2
    # empty problem
    assert lib.ql_rmsnorm_quant(fake, 128, None, None, fake, f32, 0, 128, 0, fake, fake, None, None) == 0


# ------------------------------------------------------------------ GPU

@pytest.fixture(scope="module")
def ext():
    assert torch.cuda.is_available(), "gpu tests need a CUDA device"
    from nanovllm import _ext
    assert _ext.load().ql_device_check() == 0
    return _ext


@pytest.mark.gpu
@pytest.mark.parametrize("dn", ["i8", "f8"])
@pytest.mark.parametrize("tag", ["a", "b", "c"])
def test_rmsnorm_quant_golden(ext, pgold, tag, dn):
    x, r, w = (bf(pgold[f"rms_{tag}_{k}"]).cuda() for k in ("x", "r", "w"))
    q, s, h = ext.rmsnorm_quant(x, w, 1e-6, None, DT[dn], want_bf16=True)
    assert_bf16_close(orc.t2n(h), pgold[f"rms_{tag}_h"], "rmsnorm h", RMS_TOL_ULP)
    q2, s2 = ext.quant_rows(h, DT[dn])
    assert torch.equal(q.view(torch.uint8), q2.view(torch.uint8)) and torch.equal(s, s2)
    res = r.clone()
    q, s, h = ext.rmsnorm_quant(x, w, 1e-6, res, DT[dn], want_bf16=True)
    assert np.array_equal(orc.t2n(res), pgold[f"rms_{tag}_res"])
    assert_bf16_close(orc.t2n(h), pgold[f"rms_{tag}_hres"], "add+rmsnorm h", RMS_TOL_ULP)
    q2, s2 = ext.quant_rows(h, DT[dn])
    assert torch.equal(q.view(torch.uint8), q2.view(torch.uint8)) and torch.equal(s, s2)
    # q-only call (no bf16 output) gives the same codes
    q3, s3, h3 = ext.rmsnorm_quant(x, w, 1e-6, r.clone(), DT[dn])
    assert h3 is None and torch.equal(q3.view(torch.uint8), q.view(torch.uint8)) and torch.equal(s3, s)


@pytest.mark.gpu
@pytest.mark.parametrize("dn", ["i8", "f8"])
@pytest.mark.parametrize("tag", ["a", "b", "c"])
def test_silu_mul_quant_golden(ext, pgold, tag, dn):
    gu = bf(pgold[f"silu_{tag}_x"]).cuda()
    q, s, h = ext.silu_mul_quant(gu, DT[dn], want_bf16=True)
    assert_bf16_close(orc.t2n(h), pgold[f"silu_{tag}_h"], "silu*mul h")
    q2, s2 = ext.quant_rows(h, DT[dn])
    assert torch.equal(q.view(torch.uint8), q2.view(torch.uint8)) and torch.equal(s, s2)
    q3, s3, h3 = ext.silu_mul_quant(gu, DT[dn])
    assert h3 is None and torch.equal(q3.view(torch.uint8), q.view(torch.uint8)) and torch.equal(s3, s)


@pytest.mark.gpu
@pytest.mark.parametrize("M,K", [(1, 1024), (16, 2048), (333, 4096), (4096, 1024), (64, 5120), (8, 8192), (5, 1536), (3, 16384), (6, 136)])
def test_rmsnorm_quant_vs_oracle(ext, M, K):
    g = torch.Generator().manual_seed(M * 131 + K)
    x = (torch.randn(M, K, generator=g) * 3).to(torch.bfloat16)
    r = torch.randn(M, K, generator=g).to(torch.bfloat16)
    w = (1 + 0.2 * torch.randn(K, generator=g)).to(torch.bfloat16)
    ho, reso = orc.ref_rmsnorm(x, w, 1e-6, r)
    res = r.cuda()
    q, s, h = ext.rmsnorm_quant(x.cuda(), w.cuda(), 1e-6, res, torch.int8, want_bf16=True)
    assert torch.equal(res.cpu(), reso)
    assert_bf16_close(orc.t2n(h), orc.t2n(ho), f"add+rmsnorm {M}x{K}", RMS_TOL_ULP)
    qo, so = orc.quant_rows(orc.t2n(h), orc.INT8)          # pinned quantiser on the kernel's own h
    assert np.array_equal(orc.t2n(q), qo) and np.array_equal(orc.t2n(s), so)
    # against the oracle end to end: codes within one step, scales within bf16 rounding of amax
    qo2, so2 = orc.quant_rows(orc.t2n(ho), orc.INT8)
    dq = np.abs(orc.t2n(q).view(np.int8).astype(np.int32) - qo2.view(np.int8).astype(np.int32))
    assert dq.max() <= 2 and (dq > 0).mean() <= 5e-3
    assert np.allclose(orc.t2n(s), so2, rtol=2.0 ** -7)


@pytest.mark.gpu
@pytest.mark.parametrize("M,I", [(1, 3072), (16, 3072), (257, 6144), (4096, 3072), (32, 12288), (5, 1536), (7, 768), (3, 17408), (9, 8192), (2, 40)])
def test_silu_mul_quant_vs_oracle(ext, M, I):
    g = torch.Generator().manual_seed(M * 17 + I)
    gu = (torch.randn(M, 2 * I, generator=g) * 2).to(torch.bfloat16)
    ho = orc.ref_silu_and_mul(gu)
    for dn, fmt in (("i8", orc.INT8), ("f8", orc.FP8)):
        q, s, h = ext.silu_mul_quant(gu.cuda(), DT[dn], want_bf16=True)
        assert_bf16_close(orc.t2n(h), orc.t2n(ho), f"silu*mul {M}x{I}")
        qo, so = orc.quant_rows(orc.t2n(h), fmt)
        assert np.array_equal(orc.t2n(q), qo) and np.array_equal(orc.t2n(s), so)


@pytest.mark.gpu
def test_strided_gate_up_and_layers(ext):
    """gate_up as a column slice of a wider buffer (ld > 2I) and the module mirrors chained into a
    quantized linear: fused path == unfused path bit for bit."""
    from nanovllm.layers.activation import SiluAndMul
    from nanovllm.layers.layernorm import QuantizedActivation, RMSNorm
    from nanovllm.layers.linear import MergedColumnParallelLinear, RowParallelLinear
    torch.manual_seed(3)
    M, H, I = 48, 1024, 3072
    prev = torch.get_default_dtype()
    torch.set_default_dtype(torch.bfloat16)
    try:
        with torch.device("cuda"):
            norm = RMSNorm(H)
            gate_up = MergedColumnParallelLinear(H, [I, I])
            down = RowParallelLinear(I, H)
            norm.weight.data.normal_(1.0, 0.1)
            for m in (gate_up, down):
                m.weight.data.normal_(0, 0.02)
                m.quantize_(torch.int8)
    finally:
        torch.set_default_dtype(prev)
    x = torch.randn(M, H, device="cuda").to(torch.bfloat16)
    r = torch.randn(M, H, device="cuda").to(torch.bfloat16)
    act = SiluAndMul()
    # unfused: bf16 between the ops (the reference's data flow)
    h, res = norm(x, r)
    y_ref = down(act(gate_up(h)))
    # fused: quantised activations handed from producer to linear
    qa, res2 = norm.forward_quant(x, r, torch.int8)
    assert isinstance(qa, QuantizedActivation) and torch.equal(res, res2)
    y = down(act.forward_quant(gate_up(qa), torch.int8))
    assert torch.equal(y, y_ref)
    wide = torch.randn(M, 2 * I + 64, device="cuda").to(torch.bfloat16)
    q1, s1, _ = ext.silu_mul_quant(wide[:, :2 * I], torch.int8)
    q2, s2, _ = ext.silu_mul_quant(wide[:, :2 * I].contiguous(), torch.int8)
    assert torch.equal(q1, q2) and torch.equal(s1, s2)


@pytest.mark.gpu
def test_producers_empty_batch(ext):
    """M = 0 (an empty scheduler step): nothing is launched, shapes are kept."""
    n0 = ext.launch_count()
    w = torch.ones(1024, device="cuda", dtype=torch.bfloat16)
    q, s, h = ext.rmsnorm_quant(torch.empty(0, 1024, device="cuda", dtype=torch.bfloat16), w, 1e-6, None,
                                torch.int8, want_bf16=True)
    assert q.shape == (0, 1024) and s.shape == (0,) and h.shape == (0, 1024)
    q, s, _ = ext.silu_mul_quant(torch.empty(0, 6144, device="cuda", dtype=torch.bfloat16), torch.float8_e4m3fn)
    assert q.shape == (0, 3072) and s.shape == (0,)
    assert ext.launch_count() == n0
