"""GPU parity tests proper: the sm_100a CUDA path (through the C ABI via nanovllm._ext)
against the CPU oracle (oracle/, test infrastructure) and the committed reference goldens.

Bars:
  L0 quantisers .......... bit-exact q and s
  L1 INT8 accumulators ... bit-exact int32 (ql_gemm_i8_acc)
  L2 scaled bf16 output .. INT8 per-row / per-group GEMM: bit-exact vs the oracle;
                           FP8 (fp32 accumulation order differs) and the split-K GEMV in
                           per-group mode: |dy| <= FP_TOL_ULP bf16 ulp of max|y|, rel-L2 <= FP_TOL_L2
  L3 reference thresholds  max|y - x@w.T| / max|x@w.T| < 0.08 (per-row, quantization.py:549-553)
                           and < 0.1 (per-group / per-tensor, :583, :615)
"""
import os

import numpy as np
import pytest
import torch

from oracle import oracle as orc

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

FP_TOL_ULP = 1.0      # of max|y| (one bf16 ulp at the largest magnitude = 2^-7 relative)
FP_TOL_L2 = 2e-3

DT = {"i8": torch.int8, "f8": torch.float8_e4m3fn}
FMT = {"i8": orc.INT8, "f8": orc.FP8}


@pytest.fixture(scope="module")
def ext():
    assert torch.cuda.is_available(), "gpu tests need a CUDA device"
    from nanovllm import _ext
    assert _ext.load().ql_device_check() == 0, "device is not sm_100"
    return _ext


def dev(a: np.ndarray, dtype) -> torch.Tensor:
    return orc.n2t(a, dtype).cuda()


def bits(t: torch.Tensor) -> np.ndarray:
    return orc.t2n(t)


def rand_bf16(shape, seed, scale=1.0):
    g = torch.Generator().manual_seed(seed)
    return (torch.randn(*shape, generator=g) * scale).to(torch.bfloat16)


def bf16_ulp_diff(a_u16, b_u16):
    def key(u):
        u = u.astype(np.int32)
        return np.where(u & 0x8000, 0x8000 - u, u)
    return np.abs(key(a_u16) - key(b_u16))


def assert_close_fp(y_u16, ref_u16, what=""):
    y, r = orc.bf16_decode(y_u16).astype(np.float64), orc.bf16_decode(ref_u16).astype(np.float64)
    amax = np.abs(r).max()
    if amax == 0:
        assert np.abs(y).max() == 0
        return
    ulp_at_max = 2.0 ** (np.floor(np.log2(amax)) - 7)
    err = np.abs(y - r).max()
    assert err <= FP_TOL_ULP * ulp_at_max, f"{what}: max err {err} > {FP_TOL_ULP} ulp(max)={ulp_at_max}"
    l2 = np.sqrt(((y - r) ** 2).sum() / max((r ** 2).sum(), 1e-30))
    assert l2 <= FP_TOL_L2, f"{what}: rel-L2 {l2}"


# ------------------------------------------------------------------ L0: quantisers

@pytest.mark.parametrize("dn", ["i8", "f8"])
@pytest.mark.parametrize("tag", ["a", "b", "c"])
def test_quant_rows_golden(ext, golden, dn, tag):
    for src, qk, sk in (("x", "qx", "sx"), ("w", "qw", "sw")):
        x = golden[f"pr_{tag}_{src}"]
        q, s = ext.quant_rows(dev(x, torch.bfloat16), DT[dn])
        qo, so = orc.quant_rows(x, FMT[dn])
        assert np.array_equal(bits(s), so)
        assert np.array_equal(bits(q), qo)
        assert np.array_equal(bits(s), golden[f"pr_{tag}_{dn}_{sk}"])
        if dn == "i8":      # the interpreter's fp8 cast is not the hardware cast (see test_oracle_golden)
            assert np.array_equal(bits(q), golden[f"pr_{tag}_{dn}_{qk}"])


@pytest.mark.parametrize("dn", ["i8", "f8"])
@pytest.mark.parametrize("M,K", [(1, 1024), (7, 2048), (33, 3072), (16, 4096), (5, 12288),
                                 (3, 136), (2, 100), (4096, 1024), (1, 8)])
def test_quant_rows_shapes(ext, dn, M, K):
    x = rand_bf16((M, K), 100 + M + K)
    if M > 2:
        x[1].zero_()
        x[2] *= 1e-3
        x[0, 0] = 3.0e4
    q, s = ext.quant_rows(x.cuda(), DT[dn])
    qo, so = orc.quant_rows(bits(x), FMT[dn])
    assert np.array_equal(bits(s), so)
    assert np.array_equal(bits(q), qo)


@pytest.mark.parametrize("dn", ["i8", "f8"])
def test_quant_rows_strided_and_empty(ext, dn):
    big = rand_bf16((9, 1024 + 64), 5).cuda()
    x = big[:, 32:32 + 1024]                      # row stride 1088, 64-byte offset
    q, s = ext.quant_rows(x, DT[dn])
    qo, so = orc.quant_rows(bits(x.contiguous()), FMT[dn])
    assert np.array_equal(bits(q), qo) and np.array_equal(bits(s), so)
    x2 = big[:, 3:3 + 1024]                       # misaligned -> scalar kernel
    q, s = ext.quant_rows(x2, DT[dn])
    qo, so = orc.quant_rows(bits(x2.contiguous()), FMT[dn])
    assert np.array_equal(bits(q), qo) and np.array_equal(bits(s), so)
    q, s = ext.quant_rows(torch.empty((0, 256), dtype=torch.bfloat16, device="cuda"), DT[dn])
    assert q.shape == (0, 256) and s.shape == (0,)


@pytest.mark.parametrize("dn", ["i8", "f8"])
@pytest.mark.parametrize("tag", ["a", "b"])
@pytest.mark.parametrize("gs", [64, 128, "K"])
def test_quant_groups_golden(ext, golden, dn, tag, gs):
    w = golden[f"pg_{tag}_w"]
    g = w.shape[1] if gs == "K" else gs
    q, s = ext.quant_groups(dev(w, torch.bfloat16), g, DT[dn])
    assert np.array_equal(bits(q), golden[f"pg_{tag}_{dn}_g{g}_q"])
    assert np.array_equal(bits(s), golden[f"pg_{tag}_{dn}_g{g}_s"])


@pytest.mark.parametrize("dn", ["i8", "f8"])
@pytest.mark.parametrize("N,K,g", [(1024, 3072, 128), (96, 2048, 256), (40, 1024, 32), (7, 512, 512)])
def test_quant_groups_shapes(ext, dn, N, K, g):
    w = rand_bf16((N, K), N + K + g, 0.02)
    w[N // 2].zero_()
    q, s = ext.quant_groups(w.cuda(), g, DT[dn])
    qo, so = orc.quant_groups(bits(w), g, FMT[dn])
    assert np.array_equal(bits(s), so) and np.array_equal(bits(q), qo)


@pytest.mark.parametrize("dn", ["i8", "f8"])
def test_quant_tensor_golden(ext, golden, dn):
    q, sc = ext.quant_tensor(dev(golden["pt_w"], torch.bfloat16), DT[dn])
    assert sc.dtype == torch.bfloat16 and sc.dim() == 0
    assert float(sc) == float(golden[f"pt_{dn}_s"][0])
    assert np.array_equal(bits(q), golden[f"pt_{dn}_q"])


# ------------------------------------------------------------------ L1: int32 accumulators

def rand_i8(shape, seed):
    g = torch.Generator().manual_seed(seed)
    return torch.randint(-127, 128, shape, generator=g, dtype=torch.int8)


@pytest.mark.parametrize("M,N,K", [
    (1, 8, 128), (5, 136, 256), (17, 24, 384), (32, 128, 128), (33, 1024, 1024), (64, 256, 512),
    (100, 130, 1152), (128, 128, 2048), (130, 24, 128), (256, 512, 1024), (300, 1000, 640),
    (1000, 136, 256), (513, 4096, 1024)])
def test_gemm_i8_acc_bit_exact(ext, M, N, K):
    qx, qw = rand_i8((M, K), M * 7 + K), rand_i8((N, K), N * 3 + K)
    if M > 1:
        qx[-1] = 127
        qw[-1] = -127           # extreme accumulators: K * 127^2
    acc = ext.gemm_i8_acc(qx.cuda(), qw.cuda())
    ref = orc.gemm_i8_acc(qx.numpy(), qw.numpy())
    assert np.array_equal(acc.cpu().numpy(), ref)


# ------------------------------------------------------------------ L2: scaled outputs

def test_gemm_rows_golden_long_k_int8(ext, golden):
    """Reference fixture d (K = 2048, |acc| > 2^24): every kernel family must reproduce the
    reference's int32 -> fp32 rounding, hence its bf16 output, through the public entry points."""
    p = "pr_d_i8_"
    qx, qw, sx, sw = golden[p + "qx"], golden[p + "qw"], golden[p + "sx"], golden[p + "sw"]
    yo, yf = orc.gemm_rows(qx, sx, qw, sw, None, orc.INT8, return_f32=True)
    y = ext.gemm_w8a8(dev(qx, torch.int8), dev(sx, torch.float32), dev(qw, torch.int8),
                      dev(sw, torch.float32), None, torch.int8)
    assert np.array_equal(bits(y), yo)
    assert bf16_ulp_diff(bits(y), golden[p + "y"]).max() <= 1      # interpreter truncates the last cast
    y2 = ext.linear(dev(golden["pr_d_x"], torch.bfloat16), dev(qw, torch.int8), dev(sw, torch.float32),
                    None, torch.int8)                              # quantiser + decode kernel
    assert np.array_equal(bits(y2), yo)
    acc = ext.gemm_i8_acc(dev(qx, torch.int8), dev(qw, torch.int8)).cpu().numpy()
    assert np.array_equal(acc, orc.gemm_i8_acc(qx, qw)) and np.abs(acc).max() > (1 << 24)
    # the same operands, repeated to a prefill-sized batch, through the CTA-pair kernel
    rep = 64
    yb = ext.gemm_w8a8(dev(np.tile(qx, (rep, 1)), torch.int8), dev(np.tile(sx, rep), torch.float32),
                       dev(qw, torch.int8), dev(sw, torch.float32), None, torch.int8)
    assert np.array_equal(bits(yb), np.tile(yo, (rep, 1)))


@pytest.mark.parametrize("tag", ["a", "b", "c"])
@pytest.mark.parametrize("dn", ["i8", "f8"])
def test_gemm_rows_golden(ext, golden, tag, dn):
    """Reference Triton fixtures (TRITON_INTERPRET) -> tcgen05 GEMM on their qx/qw/sx/sw."""
    p = f"pr_{tag}_{dn}_"
    qx, qw = golden[p + "qx"], golden[p + "qw"]
    K = qx.shape[1]
    y = ext.gemm_w8a8(dev(qx, DT[dn]), dev(golden[p + "sx"], torch.float32), dev(qw, DT[dn]),
                      dev(golden[p + "sw"], torch.float32), None, DT[dn])
    yo, yf = orc.gemm_rows(qx, golden[p + "sx"], qw, golden[p + "sw"], None, FMT[dn], return_f32=True)
    if dn == "i8":
        assert np.array_equal(bits(y), yo)
        # the interpreter truncates fp32->bf16; hardware rounds to nearest even (<= 1 ulp apart)
        assert bf16_ulp_diff(bits(y), golden[p + "y"]).max() <= 1
    else:
        assert_close_fp(bits(y), yo, p)
    assert K % 128 == 0


@pytest.mark.parametrize("dn", ["i8", "f8"])
@pytest.mark.parametrize("bias", [False, True])
@pytest.mark.parametrize("M,N,K", [(17, 136, 256), (64, 1024, 1024), (128, 4096, 1024),
                                   (200, 1024, 2048), (257, 520, 384), (1024, 256, 128)])
def test_gemm_rows_vs_oracle(ext, dn, bias, M, N, K):
    x, w = rand_bf16((M, K), M + N), rand_bf16((N, K), K + N, 0.02)
    b = rand_bf16((N,), 9) if bias else None
    qx, sx = orc.quant_rows(bits(x), FMT[dn])
    qw, sw = orc.quant_rows(bits(w), FMT[dn])
    y = ext.gemm_w8a8(dev(qx, DT[dn]), dev(sx, torch.float32), dev(qw, DT[dn]), dev(sw, torch.float32),
                      None if b is None else b.cuda(), DT[dn])
    yo = orc.gemm_rows(qx, sx, qw, sw, None if b is None else bits(b), FMT[dn])
    if dn == "i8":
        assert np.array_equal(bits(y), yo)
    else:
        assert_close_fp(bits(y), yo)


@pytest.mark.parametrize("dn", ["i8", "f8"])
@pytest.mark.parametrize("g", [128, 256])
@pytest.mark.parametrize("M,N,K", [(17, 136, 256), (64, 512, 1024), (128, 1024, 3072), (300, 264, 512)])
def test_gemm_groups_vs_oracle(ext, dn, g, M, N, K):
    x, w = rand_bf16((M, K), M + N + g), rand_bf16((N, K), K + N + g, 0.02)
    w[3, :g] *= 30.0
    b = rand_bf16((N,), 11)
    qx, sx = orc.quant_rows(bits(x), FMT[dn])
    qw, sw = orc.quant_groups(bits(w), g, FMT[dn])
    y = ext.gemm_w8a8(dev(qx, DT[dn]), dev(sx, torch.float32), dev(qw, DT[dn]), dev(sw, torch.float32),
                      b.cuda(), DT[dn], g)
    yo = orc.gemm_groups(qx, sx, qw, sw, bits(b), g, FMT[dn])
    if dn == "i8" and not ext.linear_path(M, N, K, g).startswith("skinny"):
        assert np.array_equal(bits(y), yo)
    else:       # fp32 group sums: the cluster split-K kernel adds them in a different (fixed) order
        assert_close_fp(bits(y), yo)


@pytest.mark.parametrize("dn", ["i8", "f8"])
@pytest.mark.parametrize("M,N,K,g", [(512, 4096, 1024, 128), (4096, 1024, 3072, 128), (4096, 6144, 1024, 128),
                                     (1000, 520, 2048, 256)])
def test_gemm_groups_prefill_sizes_vs_oracle(ext, dn, M, N, K, g):
    """Per-group at the sizes it is benchmarked at (M >= 512: the tiled per-group GEMM, BT = 128/256
    token tiles, 8 epilogue warps).  INT8: bit-exact (exact int32 group partials, the oracle's fmaf
    promotion order); FP8: the stated tolerance."""
    x, w = rand_bf16((M, K), M + N + g + 1), rand_bf16((N, K), K + N + g + 1, 0.02)
    w[3, :g] *= 30.0
    w[N - 1, K - g:] *= 0.01
    b = rand_bf16((N,), 31)
    qx, sx = orc.quant_rows(bits(x), FMT[dn])
    qw, sw = orc.quant_groups(bits(w), g, FMT[dn])
    assert not ext.linear_path(M, N, K, g).startswith("skinny")
    y = ext.gemm_w8a8(dev(qx, DT[dn]), dev(sx, torch.float32), dev(qw, DT[dn]), dev(sw, torch.float32),
                      b.cuda(), DT[dn], g)
    yo = orc.gemm_groups(qx, sx, qw, sw, bits(b), g, FMT[dn])
    if dn == "i8":
        assert np.array_equal(bits(y), yo)
    else:
        assert_close_fp(bits(y), yo)


@pytest.mark.parametrize("dn", ["i8", "f8"])
@pytest.mark.parametrize("M,N,K,g,bias", [(1030, 3330, 2048, 128, True), (2100, 1800, 2048, 512, False),
                                          (1100, 3080, 3072, 256, True), (2048, 6144, 4096, 128, False)])
def test_gemm_groups_pair_kernel_vs_oracle(ext, dn, M, N, K, g, bias):
    """gemm2g_kernel (CTA pair, 16 epilogue warps, running sums in registers), dispatched from 64 pair
    tiles and K >= 2048: M and N tails, bias, groups of 128 / 256 / 512 (every 128-byte K block is
    promoted on its own, as in the oracle), up to 32 scale groups (4 panels).  INT8 bit-exact."""
    assert ext.linear_path(M, N, K, g) == "gemm_tcgen05_2cta_grouped"
    x, w = rand_bf16((M, K), M + N + g + 2), rand_bf16((N, K), K + N + g + 2, 0.02)
    w[5, g:2 * g] *= 40.0
    b = rand_bf16((N,), 41) if bias else None
    qx, sx = orc.quant_rows(bits(x), FMT[dn])
    qw, sw = orc.quant_groups(bits(w), g, FMT[dn])
    args = (dev(qx, DT[dn]), dev(sx, torch.float32), dev(qw, DT[dn]), dev(sw, torch.float32),
            None if b is None else b.cuda(), DT[dn], g)
    y = ext.gemm_w8a8(*args)
    yo = orc.gemm_groups(qx, sx, qw, sw, None if b is None else bits(b), g, FMT[dn])
    if dn == "i8":
        assert np.array_equal(bits(y), yo)
    else:
        assert_close_fp(bits(y), yo)


def test_gemm_groups_pair_equals_single_cta(ext, monkeypatch):
    """Same arithmetic in both per-group GEMM kernels: identical bits, FP8 included."""
    import subprocess, sys, textwrap
    code = textwrap.dedent("""
        import sys, torch
        sys.path[:0] = [%r, %r]
        from nanovllm import _ext
        g = torch.Generator(device="cuda").manual_seed(5)
        M, N, K, gs = 520, 1032, 1024, 128
        x = torch.randn(M, K, generator=g, device="cuda").to(torch.bfloat16)
        w = (torch.randn(N, K, generator=g, device="cuda") * 0.02).to(torch.bfloat16)
        for dt in (torch.int8, torch.float8_e4m3fn):
            qw, sw = _ext.quant_groups(w, gs, dt)
            y = _ext.linear(x, qw, sw, None, dt, gs)
            torch.save(y.cpu(), sys.argv[1] + str(dt) + ".pt")
        print(_ext.linear_path(M, N, K, gs))
    """) % (ROOT, os.path.join(ROOT, "nano-vllm-hw3_b200"))
    import tempfile
    outs = {}
    with tempfile.TemporaryDirectory() as td:
        for mode in ("2", "0"):
            env = dict(os.environ, QL_GROUPED_PAIR=mode)
            r = subprocess.run([sys.executable, "-c", code, os.path.join(td, mode + "_")], env=env,
                               capture_output=True, text=True, timeout=300)
            assert r.returncode == 0, r.stderr[-2000:]
            assert r.stdout.strip() == ("gemm_tcgen05_2cta_grouped" if mode == "2" else "gemm_tcgen05")
            outs[mode] = [torch.load(os.path.join(td, mode + "_" + str(dt) + ".pt"))
                          for dt in (torch.int8, torch.float8_e4m3fn)]
    for a, b in zip(outs["2"], outs["0"]):
        assert torch.equal(a, b)


def test_gemm_rows_fp8_full_size_vs_oracle(ext):
    """FP8 per-row at a BASELINE prefill size (4096 x 4096 x 1024) against the oracle itself, not
    only the reference's 8 % threshold."""
    M, N, K = 4096, 4096, 1024
    x, w = rand_bf16((M, K), 77), rand_bf16((N, K), 78, 0.02)
    qx, sx = orc.quant_rows(bits(x), orc.FP8)
    qw, sw = orc.quant_rows(bits(w), orc.FP8)
    y = ext.gemm_w8a8(dev(qx, DT["f8"]), dev(sx, torch.float32), dev(qw, DT["f8"]), dev(sw, torch.float32),
                      None, DT["f8"])
    assert_close_fp(bits(y), orc.gemm_rows(qx, sx, qw, sw, None, orc.FP8))


# ------------------------------------------------------------------ prefill regime: CTA-pair kernel

PAIR_SHAPES = [(512, 4096, 128), (1024, 2048, 1024), (1000, 2072, 384), (777, 4096, 256),
               (4096, 1024, 2048), (2304, 1032, 640),
               # 128 <= M < 512: dispatched to the pair kernel since round-1n (half-empty token tiles)
               (128, 136, 256), (129, 520, 384), (200, 4096, 1024), (300, 1032, 640), (255, 256, 128)]


@pytest.mark.parametrize("M,N,K", PAIR_SHAPES)
def test_gemm_pair_i8_acc_bit_exact(ext, M, N, K):
    """cta_group::2 kernel (256x256 pair tiles): raw int32 accumulators, tails in M and N."""
    assert ext.linear_path(M, N, K) == "gemm_tcgen05_2cta"
    qx, qw = rand_i8((M, K), M * 5 + K), rand_i8((N, K), N * 11 + K)
    qx[-1] = 127
    qw[-1] = -127
    qx[M // 2] = -127
    qw[N // 2 + 1] = -127
    acc = ext.gemm_i8_acc(qx.cuda(), qw.cuda())
    ref = orc.gemm_i8_acc(qx.numpy(), qw.numpy())
    assert np.array_equal(acc.cpu().numpy(), ref)


@pytest.mark.parametrize("dn", ["i8", "f8"])
@pytest.mark.parametrize("bias", [False, True])
@pytest.mark.parametrize("M,N,K", PAIR_SHAPES)
def test_gemm_pair_rows_vs_oracle(ext, dn, bias, M, N, K):
    assert ext.linear_path(M, N, K) == "gemm_tcgen05_2cta"
    x, w = rand_bf16((M, K), M + N + 3), rand_bf16((N, K), K + N + 3, 0.02)
    b = rand_bf16((N,), 21) if bias else None
    qx, sx = orc.quant_rows(bits(x), FMT[dn])
    qw, sw = orc.quant_rows(bits(w), FMT[dn])
    y = ext.gemm_w8a8(dev(qx, DT[dn]), dev(sx, torch.float32), dev(qw, DT[dn]), dev(sw, torch.float32),
                      None if b is None else b.cuda(), DT[dn])
    yo = orc.gemm_rows(qx, sx, qw, sw, None if b is None else bits(b), FMT[dn])
    if dn == "i8":
        assert np.array_equal(bits(y), yo)
    else:
        assert_close_fp(bits(y), yo)


def test_gemm_pair_strided_output_and_unaligned_fallback(ext):
    """ql_linear with a strided y (ldy > N) goes through the TMA-store epilogue; an output whose
    pitch is not a multiple of 16 bytes falls back to the single-CTA kernel.  Both bit-exact."""
    M, N, K = 1024, 2048, 512
    x, w = rand_bf16((M, K), 5), rand_bf16((N, K), 6, 0.02)
    qw, sw = orc.quant_rows(bits(w), orc.INT8)
    qx, sx = orc.quant_rows(bits(x), orc.INT8)
    yo = orc.gemm_rows(qx, sx, qw, sw, None, orc.INT8)
    for pitch in (N + 64, N + 3):
        buf = torch.full((M, pitch), 7.0, dtype=torch.bfloat16, device="cuda")
        out = buf[:, :N]
        ext.linear(x.cuda(), dev(qw, torch.int8), dev(sw, torch.float32), None, torch.int8, 0, out=out)
        assert np.array_equal(bits(out.contiguous()), yo)
        assert bool((buf[:, N:] == 7.0).all())          # nothing written past column N


# ------------------------------------------------------------------ decode regime: fused GEMV

@pytest.mark.parametrize("dn", ["i8", "f8"])
@pytest.mark.parametrize("M", [1, 2, 5, 8, 9, 16])
@pytest.mark.parametrize("N,K", [(4096, 1024), (1024, 2048), (6144, 1024), (1024, 3072), (136, 256),
                                 (8, 128), (12, 4096)])
def test_gemv_rows(ext, dn, M, N, K):
    x, w = rand_bf16((M, K), M + N + 1), rand_bf16((N, K), K + N + 1, 0.02)
    b = rand_bf16((N,), 13) if (M % 2) else None
    qw, sw = orc.quant_rows(bits(w), FMT[dn])
    y = ext.gemv_w8a8(x.cuda(), dev(qw, DT[dn]), dev(sw, torch.float32),
                      None if b is None else b.cuda(), DT[dn])
    qx, sx = orc.quant_rows(bits(x), FMT[dn])
    yo = orc.gemm_rows(qx, sx, qw, sw, None if b is None else bits(b), FMT[dn])
    if dn == "i8":
        assert np.array_equal(bits(y), yo)          # exact int32 sums, same epilogue order
    else:
        assert_close_fp(bits(y), yo)


@pytest.mark.parametrize("dn", ["i8", "f8"])
@pytest.mark.parametrize("M", [1, 7, 16])
@pytest.mark.parametrize("N,K,g", [(4096, 1024, 128), (1024, 3072, 128), (520, 512, 256), (6144, 1024, 1024)])
def test_gemv_groups(ext, dn, M, N, K, g):
    x, w = rand_bf16((M, K), M + N + 2), rand_bf16((N, K), K + N + 2, 0.02)
    w[1, :g] *= 25.0
    qw, sw = orc.quant_groups(bits(w), g, FMT[dn])
    y = ext.gemv_w8a8(x.cuda(), dev(qw, DT[dn]), dev(sw, torch.float32), None, DT[dn], g)
    qx, sx = orc.quant_rows(bits(x), FMT[dn])
    yo = orc.gemm_groups(qx, sx, qw, sw, None, g, FMT[dn])
    assert_close_fp(bits(y), yo)                     # split-K changes the fp32 summation order


# ------------------------------------------------------------------ decode regime: cluster split-K tcgen05 kernel

def _force_split(monkeypatch, S):
    monkeypatch.setenv("QL_SKINNY_ALWAYS", "1")     # these tests target the decode kernel at every M <= 256
    if S:
        monkeypatch.setenv("QL_SKINNY_S", str(S))
    else:
        monkeypatch.delenv("QL_SKINNY_S", raising=False)


def test_skinny_dispatch(ext):
    """Dispatch measured in profiles/r1n_dispatch_crossover.log."""
    assert ext.linear_path(1, 4096, 1024) == "quant+skinny_tcgen05"
    assert ext.linear_path(16, 1024, 3072, 128) == "quant+skinny_tcgen05"
    assert ext.linear_path(16, 24576, 4096) == "quant+skinny_tcgen05"   # long K slice: two launches
    assert ext.linear_path(17, 4096, 1024) == "skinny_tcgen05"
    assert ext.linear_path(64, 4096, 1024) == "gemm_tcgen05"            # short K: no split-K gain
    assert ext.linear_path(64, 6144, 4096) == "skinny_tcgen05"
    assert ext.linear_path(128, 1024, 3072) == "skinny_tcgen05"         # K >= 2N
    assert ext.linear_path(128, 6144, 4096) == "gemm_tcgen05_2cta"
    assert ext.linear_path(256, 4096, 12288) == "skinny_tcgen05"
    assert ext.linear_path(256, 6144, 4096) == "gemm_tcgen05_2cta"
    assert ext.linear_path(128, 6144, 4096, 128) == "skinny_tcgen05"
    assert ext.linear_path(129, 6144, 4096, 128) == "gemm_tcgen05"
    assert ext.linear_path(257, 4096, 12288) == "gemm_tcgen05_2cta"


@pytest.mark.parametrize("S", [0, 1, 2, 4, 8])
@pytest.mark.parametrize("M,N,K", [(1, 128, 1024), (3, 1000, 1152), (16, 520, 2048), (17, 4096, 1024),
                                   (40, 136, 1152), (64, 1024, 3072), (100, 1000, 1152),
                                   (128, 6144, 1024), (200, 264, 1024), (256, 1024, 2048)])
def test_skinny_i8_acc_bit_exact(ext, monkeypatch, S, M, N, K):
    """Every split factor (even and uneven K slices, M / N tails): the DSMEM reduce-scatter of
    int32 partials must reproduce the exact accumulators."""
    _force_split(monkeypatch, S)
    qx, qw = rand_i8((M, K), M * 13 + K), rand_i8((N, K), N * 17 + K)
    qx[-1] = 127
    qw[-1] = -127
    qw[N // 2] = 127
    acc = ext.gemm_i8_acc(qx.cuda(), qw.cuda())
    ref = orc.gemm_i8_acc(qx.numpy(), qw.numpy())
    assert np.array_equal(acc.cpu().numpy(), ref)


@pytest.mark.parametrize("S", [0, 1, 2, 4, 8])
@pytest.mark.parametrize("dn", ["i8", "f8"])
@pytest.mark.parametrize("M,N,K", [(1, 4096, 1024), (2, 1024, 3072), (7, 1000, 1152), (16, 6144, 1024),
                                   (12, 136, 4096), (16, 1024, 8192)])
@pytest.mark.parametrize("fused", [0, 1])
def test_skinny_fused_rows(ext, monkeypatch, fused, S, dn, M, N, K):
    """Decode regime through ql_linear, strided x and y.  fused=1: quantiser inside the kernel
    (partial amax combined across the cluster, codes written in the UMMA layout by the CTA);
    fused=0 (the default dispatch): quant_rows launch + TMA-fed split-K kernel."""
    _force_split(monkeypatch, S)
    monkeypatch.setenv("QL_FUSED", str(fused))
    xb = rand_bf16((M, K + 64), M + N + 5)
    xb[0, 40] = 200.0                              # the row maximum sits in the first K slice only
    if M > 2:
        xb[2].zero_()                              # all-zero row -> amax clamp 1e-8
    x = xb[:, 32:32 + K]
    w = rand_bf16((N, K), K + N + 5, 0.02)
    b = rand_bf16((N,), 15) if (M % 2) else None
    qw, sw = orc.quant_rows(bits(w), FMT[dn])
    buf = torch.full((M, N + 8), 3.0, dtype=torch.bfloat16, device="cuda")
    out = buf[:, :N]
    ext.linear(x.cuda(), dev(qw, DT[dn]), dev(sw, torch.float32), None if b is None else b.cuda(),
               DT[dn], 0, out=out)
    qx, sx = orc.quant_rows(bits(x.contiguous()), FMT[dn])
    yo = orc.gemm_rows(qx, sx, qw, sw, None if b is None else bits(b), FMT[dn])
    if dn == "i8":
        assert np.array_equal(bits(out.contiguous()), yo)
    else:
        assert_close_fp(bits(out.contiguous()), yo)
    assert bool((buf[:, N:] == 3.0).all())


@pytest.mark.parametrize("S", [0, 1, 4])
@pytest.mark.parametrize("dn", ["i8", "f8"])
@pytest.mark.parametrize("M,N,K,g", [(1, 4096, 1024, 128), (16, 1024, 3072, 128), (5, 520, 1024, 256),
                                     (33, 1024, 1024, 128), (64, 264, 2048, 256), (128, 4096, 1024, 128)])
@pytest.mark.parametrize("fused", [0, 1])
def test_skinny_groups(ext, monkeypatch, fused, S, dn, M, N, K, g):
    _force_split(monkeypatch, S)
    monkeypatch.setenv("QL_FUSED", str(fused))
    x, w = rand_bf16((M, K), M + N + 6), rand_bf16((N, K), K + N + 6, 0.02)
    w[1, :g] *= 25.0
    b = rand_bf16((N,), 17)
    qw, sw = orc.quant_groups(bits(w), g, FMT[dn])
    y = ext.linear(x.cuda(), dev(qw, DT[dn]), dev(sw, torch.float32), b.cuda(), DT[dn], g)
    qx, sx = orc.quant_rows(bits(x), FMT[dn])
    yo = orc.gemm_groups(qx, sx, qw, sw, bits(b), g, FMT[dn])
    assert_close_fp(bits(y), yo)


@pytest.mark.parametrize("dn", ["i8", "f8"])
@pytest.mark.parametrize("M", [17, 48, 128, 130, 256])
def test_skinny_linear_rows_full_shapes(ext, dn, M):
    """ql_linear (quant_rows_kernel + skinny kernel, PDL between them) on the Qwen3 decode shapes."""
    for N, K in [(4096, 1024), (1024, 2048), (6144, 1024), (1024, 3072), (6144, 4096), (4096, 12288)]:
        x, w = rand_bf16((M, K), M + N + 7), rand_bf16((N, K), K + N + 7, 0.02)
        qw, sw = orc.quant_rows(bits(w), FMT[dn])
        y = ext.linear(x.cuda(), dev(qw, DT[dn]), dev(sw, torch.float32), None, DT[dn])
        qx, sx = orc.quant_rows(bits(x), FMT[dn])
        yo = orc.gemm_rows(qx, sx, qw, sw, None, FMT[dn])
        if dn == "i8":
            assert np.array_equal(bits(y), yo), (N, K)
        else:
            assert_close_fp(bits(y), yo, str((N, K)))


def test_skinny_back_to_back_pdl(ext):
    """A chain of dependent linears (each reads the previous output) launched back to back: with
    programmatic dependent launch the next kernel's weight prefetch overlaps the previous one,
    but its activation reads must still see the finished output."""
    torch.manual_seed(3)
    K = 1024
    ws = [(torch.randn(K, K, device="cuda") * 0.05).to(torch.bfloat16) for _ in range(6)]
    qs = [ext.quant_rows(w, torch.int8) for w in ws]
    for M in (4, 64):
        x = torch.randn(M, K, device="cuda").to(torch.bfloat16)
        h = x
        for q, s in qs:
            h = ext.linear(h, q, s, None, torch.int8)
        torch.cuda.synchronize()
        r = x
        for q, s in qs:                            # same chain, synchronising after every call
            r = ext.linear(r, q, s, None, torch.int8)
            torch.cuda.synchronize()
        assert torch.equal(h, r)


# ------------------------------------------------------------------ reference-style API tests

def _rel_err(y, y_ref):
    return float((y.float() - y_ref.float()).abs().max() / y_ref.float().abs().max())


@pytest.mark.parametrize("dtype", [torch.int8, torch.float8_e4m3fn])
@pytest.mark.parametrize("M,N,K", [(1, 1536, 1536), (10, 1536, 1536), (4096, 4096, 4096)])
def test_quant_mm(ext, dtype, M, N, K):
    """The reference's own test_quant_mm (quantization.py:527-555), seeded, with its threshold."""
    from nanovllm.utils.quantization import per_row_matmul, per_row_quant
    torch.manual_seed(M + N)
    x = torch.randn(M, K, dtype=torch.bfloat16, device="cuda")
    w = torch.randn(N, K, dtype=torch.bfloat16, device="cuda")
    b = torch.randn(N, dtype=torch.bfloat16, device="cuda")
    qw, sw = per_row_quant(w, dtype)
    assert qw.dtype == dtype and sw.dtype == torch.float32 and sw.shape == (N,)
    y = per_row_matmul(x, qw, sw, b, dtype)
    y_ref = x @ w.T + b
    assert y.shape == y_ref.shape and y.dtype == torch.bfloat16
    assert _rel_err(y, y_ref) < 0.08
    y3 = per_row_matmul(x.view(1, M, K), qw, sw, None, dtype)       # 3-D input (test_ppl.py shape)
    assert y3.shape == (1, M, N)


@pytest.mark.parametrize("dtype", [torch.int8, torch.float8_e4m3fn])
@pytest.mark.parametrize("g", [128, 256])
@pytest.mark.parametrize("M,N,K", [(1, 512, 512), (10, 1024, 1024), (100, 1024, 1024)])
def test_group_quant_mm(ext, dtype, g, M, N, K):
    """test_group_quant_mm (quantization.py:588-619) against bf16 matmul AND the reference's
    weight-only dequant+matmul (oracle restatement), threshold < 0.1 as the reference."""
    from nanovllm.utils.quantization import per_group_matmul, per_group_quant
    torch.manual_seed(M + N + g)
    x = torch.randn(M, K, dtype=torch.bfloat16, device="cuda")
    w = torch.randn(N, K, dtype=torch.bfloat16, device="cuda")
    b = torch.randn(N, dtype=torch.bfloat16, device="cuda")
    qw, sw = per_group_quant(w, g, dtype)
    assert sw.shape == (N, K // g)
    y = per_group_matmul(x, qw, sw, b, g, dtype)
    assert _rel_err(y, x @ w.T + b) < 0.1
    y_ref = orc.ref_per_group_matmul(x.cpu(), qw.cpu(), sw.cpu(), b.cpu(), g, dtype)
    assert _rel_err(y.cpu(), y_ref) < 0.1


@pytest.mark.parametrize("dtype", [torch.int8, torch.float8_e4m3fn])
@pytest.mark.parametrize("M,N,K", [(1, 512, 512), (10, 1024, 1024), (32, 2048, 2048)])
def test_tensor_quant_mm(ext, dtype, M, N, K):
    """test_tensor_quant_mm (quantization.py:558-585), threshold < 0.1."""
    from nanovllm.utils.quantization import per_tensor_matmul, per_tensor_quant
    torch.manual_seed(M + N)
    x = torch.randn(M, K, dtype=torch.bfloat16, device="cuda")
    w = torch.randn(N, K, dtype=torch.bfloat16, device="cuda")
    b = torch.randn(N, dtype=torch.bfloat16, device="cuda")
    qw, sw = per_tensor_quant(w, dtype)
    y = per_tensor_matmul(x, qw, sw, b, dtype)
    assert _rel_err(y, x @ w.T + b) < 0.1


def test_assertions_like_reference(ext):
    from nanovllm.utils.quantization import per_group_quant, per_row_matmul, per_row_quant
    x = torch.randn(4, 256, dtype=torch.bfloat16, device="cuda")
    with pytest.raises(AssertionError):
        per_row_quant(x, torch.float16)                       # dtype not in [int8, e4m3]
    with pytest.raises(AssertionError):
        per_row_quant(x.view(2, 2, 256), torch.int8)          # x.dim() != 2
    with pytest.raises(AssertionError):
        per_group_quant(x, 96, torch.int8)                    # K % group_size != 0
    w = torch.randn(8, 192, dtype=torch.bfloat16, device="cuda")
    qw, sw = per_row_quant(w, torch.int8)
    with pytest.raises(AssertionError):
        per_row_matmul(torch.randn(4, 192, dtype=torch.bfloat16, device="cuda"), qw, sw, None,
                       torch.int8)                            # K % 128 != 0


# ------------------------------------------------------------------ full-size properties

@pytest.mark.parametrize("M,N,K", [(4096, 4096, 1024), (8192, 1024, 3072), (2048, 6144, 4096)])
def test_full_size_int8_checksums(ext, M, N, K):
    """BASELINE-size shapes: row and column checksums of the int32 accumulators, computed
    independently in int64 on the device (sum_n acc[m,n] == qx[m,:] . sum_n qw[n,:])."""
    qx, qw = rand_i8((M, K), 1).cuda(), rand_i8((N, K), 2).cuda()
    acc = ext.gemm_i8_acc(qx, qw).to(torch.int64)
    col_w = qw.to(torch.int64).sum(0)
    row_x = qx.to(torch.int64).sum(0)
    assert torch.equal(acc.sum(1), (qx.to(torch.int64) * col_w).sum(1))
    assert torch.equal(acc.sum(0), (qw.to(torch.int64) * row_x).sum(1))
    # spot-check 64 random entries exactly
    g = torch.Generator().manual_seed(3)
    mi = torch.randint(0, M, (64,), generator=g).cuda()
    ni = torch.randint(0, N, (64,), generator=g).cuda()
    exact = (qx[mi].to(torch.int64) * qw[ni].to(torch.int64)).sum(1)
    assert torch.equal(acc[mi, ni], exact)


@pytest.mark.parametrize("dn", ["i8", "f8"])
def test_full_size_linearity_and_path_agreement(ext, dn):
    """Scaling an activation row by 2 (exact in bf16) must scale that output row by exactly 2
    (per-token scales absorb it), and the GEMV and GEMM paths must agree on shared rows."""
    M, N, K = 4096, 4096, 1024
    x = rand_bf16((M, K), 21).cuda()
    w = rand_bf16((N, K), 22, 0.02).cuda()
    qw, sw = ext.quant_rows(w, DT[dn])
    y1 = ext.linear(x, qw, sw, None, DT[dn])
    y2 = ext.linear(x * 2, qw, sw, None, DT[dn])
    assert torch.equal(y2.float(), y1.float() * 2)
    yv = ext.linear(x[:16].contiguous(), qw, sw, None, DT[dn])      # M <= 16 -> GEMV path
    if dn == "i8":
        assert torch.equal(yv, y1[:16])
    else:
        assert_close_fp(bits(yv), bits(y1[:16].contiguous()))


# ------------------------------------------------------------------ layers, graphs

def test_linear_dispatch_and_cuda_graph(ext):
    from nanovllm.layers.linear import MergedColumnParallelLinear, RowParallelLinear
    torch.manual_seed(0)
    with torch.device("cuda"):
        prev = torch.get_default_dtype()
        torch.set_default_dtype(torch.bfloat16)
        try:
            up = MergedColumnParallelLinear(1024, [3072, 3072])
            down = RowParallelLinear(3072, 1024)
        finally:
            torch.set_default_dtype(prev)
    up.weight.data.normal_(0, 0.02)
    down.weight.data.normal_(0, 0.02)
    w_up = up.weight.data.clone()
    up.quantize_(torch.int8)
    down.quantize_(torch.float8_e4m3fn, "per_group", 128)
    assert up.weight is None and up.qw.dtype == torch.int8 and up.sw.shape == (6144,)
    assert down.sw.shape == (1024, 24)
    for M in (4, 48):
        x = torch.randn(M, 1024, dtype=torch.bfloat16, device="cuda")
        h = up(x)
        qx, sx = orc.quant_rows(bits(x), orc.INT8)
        yo = orc.gemm_rows(qx, sx, bits(up.qw), bits(up.sw), None, orc.INT8)
        assert np.array_equal(bits(h), yo)
        assert _rel_err(h, x @ w_up.T) < 0.08
        y = down(h[:, :3072].contiguous())
        assert y.shape == (M, 1024)
    # CUDA-graph capture (the reference captures decode, model_runner.py:316-356)
    xs = torch.randn(8, 1024, dtype=torch.bfloat16, device="cuda")
    eager = up(xs).clone()
    g = torch.cuda.CUDAGraph()
    s = torch.cuda.Stream()
    s.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(s):
        up(xs)
        with torch.cuda.graph(g, stream=s):
            out = up(xs)
    torch.cuda.current_stream().wait_stream(s)
    xs.copy_(xs * 0.5)
    g.replay()
    torch.cuda.synchronize()
    assert torch.equal(out, up(xs))
    assert torch.equal(out.float() * 2, eager.float())


# ------------------------------------------------------------------ shared workspace under PDL
def test_linear_shared_workspace_back_to_back(ext):
    """nanovllm._ext keeps ONE scratch buffer per (device, stream) and the kernels are
    launched with programmatic dependent launch, so call i+1's quantiser may start while call i's
    GEMM still reads the buffer -- it must not write before that GEMM is done.  Differently sized
    linears (prefill tiles, ragged M, decode) are issued back to back with no synchronisation and
    every output must carry the bits of the same call made alone with a private workspace."""
    lib = ext.load()
    shapes = [(4096, 1024, 1024), (17, 4096, 1024), (300, 1024, 3072), (2048, 2048, 2048), (16, 1024, 4096),
              (1000, 6144, 1024), (64, 1024, 2048), (4096, 1024, 3072), (33, 512, 384), (129, 4096, 1024)]
    for kind, group in (("i8", 0), ("f8", 0), ("i8", 128)):
        dtype = DT[kind]
        probs = []
        for i, (M, N, K) in enumerate(shapes):
            g = torch.Generator(device="cuda").manual_seed(1000 + i)
            x = torch.randn(M, K, generator=g, device="cuda").to(torch.bfloat16)
            w = (torch.randn(N, K, generator=g, device="cuda") * 0.02).to(torch.bfloat16)
            qw, sw = ext.quant_groups(w, group, dtype) if group else ext.quant_rows(w, dtype)
            probs.append((x, qw, sw))
        # reference bits: each call alone, its own workspace, synchronised on both sides
        refs = []
        for x, qw, sw in probs:
            M, K = x.shape
            N = qw.shape[0]
            y = torch.empty(M, N, dtype=torch.bfloat16, device="cuda")
            nbytes = lib.ql_workspace_bytes(M, K)
            ws = torch.empty(max(nbytes, 16), dtype=torch.uint8, device="cuda")
            torch.cuda.synchronize()
            ext._check(lib.ql_linear(x.data_ptr(), K, qw.data_ptr(), sw.data_ptr(), None, y.data_ptr(), N, M, N, K,
                                     ext.fmt_of(dtype), group, ws.data_ptr(), nbytes,
                                     torch.cuda.current_stream().cuda_stream))
            torch.cuda.synchronize()
            refs.append(y)
        for rep in range(3):
            order = list(range(len(probs))) if rep == 0 else list(torch.randperm(len(probs)).tolist())
            outs = {i: ext.linear(*probs[i], None, dtype, group) for i in order for _ in (0,)}
            torch.cuda.synchronize()
            for i in order:
                assert torch.equal(outs[i], refs[i]), (kind, group, rep, shapes[i])
