"""The prefill-sized tensor-parallel collective (csrc/tp_collective.cu + the PUSH epilogue of
gemm2_kernel) on ONE GPU: W ranks emulated inside one process (``ql_ar_init_local``), so the
single-GPU test tier executes the very kernels, flags and tensor maps the multi-GPU path uses
(tests/test_tp_collective.py runs the same checks across real GPUs).

Replaces ``dist.all_reduce(y)`` (reference nanovllm/layers/linear.py:150-152) + the next linear's
``per_row_quant`` (utils/quantization.py:147-163).  Bar: bit-identical to

    y = bf16( sum over ranks r = 0..W-1 of float(partial_r), fp32, rank order )
    (q, s) = ql_quant_rows(y)

for every rank's copy of the gathered codes / scales / bf16 rows; partial_r = the rank's own
row-parallel GEMM output (checked through the PUSH kernel against ql_gemm_w8a8 on the same operands).
"""
import ctypes
import os

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

DT = {"i8": torch.int8, "f8": torch.float8_e4m3fn}
TOKENS, HIDDEN = 4608, 8192


@pytest.fixture(scope="module")
def ext():
    assert torch.cuda.is_available(), "gpu tests need a CUDA device"
    from nanovllm import _ext
    assert _ext.load().ql_device_check() == 0, "device is not sm_100"
    return _ext


class Emu:
    """W emulated ranks; buffers wrapped as torch tensors."""

    def __init__(self, ext, world):
        self.ext, self.lib, self.world = ext, ext.load(), world
        self.ch = 1 if world == 4 else 0          # the channels are independent copies: cover both
        # channel 1 needs the explicit opt-in (the hazard is two channels IN FLIGHT AT ONCE; these
        # tests use one channel at a time on one stream, which is safe)
        os.environ["QL_TP_ALLOW_CONCURRENT"] = "1" if self.ch else "0"
        self.lib.ql_ar_destroy()
        ext._check(self.lib.ql_ar_init_local(1 << 16, TOKENS, HIDDEN, world))

    def close(self):
        torch.cuda.synchronize()
        self.lib.ql_ar_destroy()

    def buf(self, rank, which, nbytes):
        from nanovllm._comm import _DevMem
        self.ext._check(self.lib.ql_tp_emu_set_rank(rank))
        ptr = self.lib.ql_tp_buffer(which, self.ch)
        assert ptr
        return torch.as_tensor(_DevMem(ptr, nbytes), device="cuda")

    def stream(self):
        return torch.cuda.current_stream().cuda_stream


@pytest.fixture(params=["ag_push", "ag_pull"])
def ag_mode(request, monkeypatch):
    """All-gather of the codes by posted stores or by peer loads after the done flags (QL_TP_AG_PULL)."""
    monkeypatch.setenv("QL_TP_AG_PULL", "1" if request.param == "ag_pull" else "0")
    return request.param


@pytest.fixture(params=[2, 4, 8])
def emu(request, ext):
    e = Emu(ext, request.param)
    yield e
    e.close()


def _partials(world, M, N, seed):
    g = torch.Generator().manual_seed(seed)
    parts = [(torch.randn(M, N, generator=g) * (1.0 + r)).to(torch.bfloat16).cuda() for r in range(world)]
    if M > 3:
        for p in parts:
            p[1].zero_()                                  # an all-zero token: amax clamp 1e-8
        parts[0][2] *= 1e-3
    return parts


def _expected(ext, parts, dtype):
    acc = parts[0].float()
    for p in parts[1:]:
        acc = acc + p.float()                             # fp32, rank order: the kernel's definition
    y = acc.to(torch.bfloat16)
    q, s = ext.quant_rows(y, dtype) if dtype is not None else (None, None)
    return y, q, s


@pytest.mark.parametrize("dn", ["i8", "f8"])
@pytest.mark.parametrize("M,N", [(512, 1024), (1000, 1024), (300, 2048), (77, 1024), (2304, 4096),
                                 (520, 8192), (4097, 136)])
def test_pull_reduce_quant_gather(ext, emu, ag_mode, dn, M, N):
    W, dtype = emu.world, DT[dn]
    parts = _partials(W, M, N, M + N + W)
    for r in range(W):
        emu.buf(r, 0, M * N * 2).view(torch.bfloat16).view(M, N).copy_(parts[r])
    y, q, s = _expected(ext, parts, dtype)
    for rep in range(2):                                  # second call: the epoch counters advanced
        ext._check(emu.lib.ql_tp_reduce_quant_gather(M, N, ext.fmt_of(dtype), 0, emu.ch, emu.stream()))
        torch.cuda.synchronize()
        for r in range(W):
            qr = emu.buf(r, 1, M * N).view(dtype).view(M, N)
            sr = emu.buf(r, 2, M * 4).view(torch.float32)
            assert torch.equal(sr, s), (rep, r)
            assert torch.equal(qr.view(torch.uint8), q.view(torch.uint8)), (rep, r)
            qr.view(torch.uint8).fill_(0x55)              # the next call must rewrite every byte
            sr.fill_(-1.0)


@pytest.mark.parametrize("M,N", [(512, 1024), (1000, 2048), (77, 1024), (600, 4096)])
def test_pull_allreduce_bf16(ext, emu, M, N):
    W = emu.world
    parts = _partials(W, M, N, 3 * M + N + W)
    for r in range(W):
        emu.buf(r, 0, M * N * 2).view(torch.bfloat16).view(M, N).copy_(parts[r])
    y, _, _ = _expected(ext, parts, None)
    out = torch.full((W, M, N), 7.0, dtype=torch.bfloat16, device="cuda")
    ext._check(emu.lib.ql_tp_allreduce_bf16(M, N, 0, out.data_ptr(), N, emu.ch, emu.stream()))
    torch.cuda.synchronize()
    for r in range(W):
        assert torch.equal(out[r], y), r


@pytest.mark.parametrize("dn", ["i8", "f8"])
@pytest.mark.parametrize("bias", [False, True])
@pytest.mark.parametrize("M,N,K", [(512, 1024, 256), (1000, 1024, 384), (300, 2048, 128), (2304, 1032, 256),
                                   (4096, 1024, 128)])
def test_push_gemm_reduce_quant_gather(ext, emu, ag_mode, dn, bias, M, N, K):
    """GEMM + reduce-scatter in one kernel (gemm2_push_kernel): every rank's row-parallel GEMM
    stores its 256-token tiles straight into the owners' receive regions; then the owners reduce."""
    W, dtype = emu.world, DT[dn]
    g = torch.Generator().manual_seed(M + N + K + W)
    parts = []
    for r in range(W):
        x = torch.randn(M, K, generator=g).to(torch.bfloat16).cuda()
        w = (torch.randn(N, K, generator=g) * 0.02).to(torch.bfloat16).cuda()
        b = torch.randn(N, generator=g).to(torch.bfloat16).cuda() if (bias and r == 0) else None   # rank 0 only
        qx, sx = ext.quant_rows(x, dtype)
        qw, sw = ext.quant_rows(w, dtype)
        parts.append(ext.gemm_w8a8(qx, sx, qw, sw, b, dtype))            # the un-fused kernel's result
        ext._check(emu.lib.ql_tp_emu_set_rank(r))
        ext._check(emu.lib.ql_tp_gemm_push(qx.data_ptr(), sx.data_ptr(), qw.data_ptr(), sw.data_ptr(),
                                           None if b is None else b.data_ptr(), M, N, K, ext.fmt_of(dtype),
                                           emu.ch, emu.stream()))
    y, q, s = _expected(ext, parts, dtype)
    ext._check(emu.lib.ql_tp_reduce_quant_gather(M, N, ext.fmt_of(dtype), 1, emu.ch, emu.stream()))
    torch.cuda.synchronize()
    for r in range(W):
        assert torch.equal(emu.buf(r, 2, M * 4).view(torch.float32), s), r
        assert torch.equal(emu.buf(r, 1, M * N), q.view(torch.uint8).view(-1)), r


def _push_round(ext, emu, dtype, M, N, K, seed, bias=False):
    """One row-parallel linear on every emulated rank (inputs made on the GPU: the 8B shards are
    large) -> the un-fused kernel's partials, with the pushed copies sitting in the owners' areas."""
    W = emu.world
    g = torch.Generator(device="cuda").manual_seed(seed)
    parts = []
    for r in range(W):
        x = torch.randn(M, K, generator=g, device="cuda").to(torch.bfloat16)
        w = (torch.randn(N, K, generator=g, device="cuda") * 0.02).to(torch.bfloat16)
        b = torch.randn(N, generator=g, device="cuda").to(torch.bfloat16) if (bias and r == 0) else None
        qx, sx = ext.quant_rows(x, dtype)
        qw, sw = ext.quant_rows(w, dtype)
        parts.append(ext.gemm_w8a8(qx, sx, qw, sw, b, dtype))
        ext._check(emu.lib.ql_tp_emu_set_rank(r))
        ext._check(emu.lib.ql_tp_gemm_push(qx.data_ptr(), sx.data_ptr(), qw.data_ptr(), sw.data_ptr(),
                                           None if b is None else b.data_ptr(), M, N, K, ext.fmt_of(dtype),
                                           emu.ch, emu.stream()))
    return parts


@pytest.mark.parametrize("dn", ["i8", "f8"])
@pytest.mark.parametrize("K", [512, 1536, 2048, 6144])
def test_push_gemm_qwen3_8b_shards(ext, emu, dn, K):
    """BASELINE configs[3]: the row-parallel shards of Qwen3-8B (hidden 4096; o_proj K = 4096/tp,
    down_proj K = 12288/tp: 512 and 1536 at TP=8, 2048 and 6144 at TP=2) through gemm2_push_kernel
    + tp_reduce_kernel, against the un-fused GEMM + "fp32 rank-order sum -> bf16 -> ql_quant_rows"."""
    W, dtype, M, N = emu.world, DT[dn], 1280, 4096
    parts = _push_round(ext, emu, dtype, M, N, K, 8000 + K + W)
    y, q, s = _expected(ext, parts, dtype)
    ext._check(emu.lib.ql_tp_reduce_quant_gather(M, N, ext.fmt_of(dtype), 1, emu.ch, emu.stream()))
    torch.cuda.synchronize()
    for r in range(W):
        assert torch.equal(emu.buf(r, 2, M * 4).view(torch.float32), s), r
        assert torch.equal(emu.buf(r, 1, M * N), q.view(torch.uint8).view(-1)), r


def test_push_into_poisoned_receive_regions(ext, emu):
    """the reduce kernel may only read rows the GEMM epilogue's bulk stores have
    completed.  Every iteration pushes NEW data into receive regions (and result buffers) that were
    overwritten with a poison pattern after the previous one; a stale or missing row cannot match."""
    W, dtype, (M, N, K) = emu.world, torch.int8, (1024, 2048, 256)
    cap = emu.lib.ql_tp_max_tokens() * emu.lib.ql_tp_max_hidden()
    for it in range(12):
        for r in range(W):
            emu.buf(r, 0, cap * 2).fill_(0x7F)            # bf16 0x7f7f = NaN: poisons any sum it enters
            emu.buf(r, 1, M * N).fill_(0x55)
            emu.buf(r, 2, M * 4).view(torch.float32).fill_(-1.0)
        parts = _push_round(ext, emu, dtype, M, N, K, 500 + 31 * it + W, bias=bool(it & 1))
        ext._check(emu.lib.ql_tp_reduce_quant_gather(M, N, ext.fmt_of(dtype), 1, emu.ch, emu.stream()))
        _, q, s = _expected(ext, parts, dtype)
        torch.cuda.synchronize()
        for r in range(W):
            assert torch.equal(emu.buf(r, 2, M * 4).view(torch.float32), s), (it, r)
            assert torch.equal(emu.buf(r, 1, M * N), q.view(torch.uint8).view(-1)), (it, r)


def test_push_allreduce_bf16_and_graph_replay(ext, emu):
    """The bf16 form after pushed partials, captured in a CUDA graph and replayed (the engine
    captures its steps, model_runner.py:316-356): the device-side call counters keep every
    replay consistent."""
    W, dtype, (M, N, K) = emu.world, torch.int8, (768, 1024, 256)
    g = torch.Generator().manual_seed(11 + W)
    ops, parts = [], []
    for r in range(W):
        x = torch.randn(M, K, generator=g).to(torch.bfloat16).cuda()
        w = (torch.randn(N, K, generator=g) * 0.02).to(torch.bfloat16).cuda()
        qx, sx = ext.quant_rows(x, dtype)
        qw, sw = ext.quant_rows(w, dtype)
        ops.append((qx, sx, qw, sw))
        parts.append(ext.gemm_w8a8(qx, sx, qw, sw, None, dtype))
    y, q, s = _expected(ext, parts, dtype)
    out = torch.zeros((W, M, N), dtype=torch.bfloat16, device="cuda")

    def one_round():
        for r, (qx, sx, qw, sw) in enumerate(ops):
            ext._check(emu.lib.ql_tp_emu_set_rank(r))
            ext._check(emu.lib.ql_tp_gemm_push(qx.data_ptr(), sx.data_ptr(), qw.data_ptr(), sw.data_ptr(), None,
                                               M, N, K, ext.fmt_of(dtype), emu.ch, emu.stream()))
        ext._check(emu.lib.ql_tp_allreduce_bf16(M, N, 1, out.data_ptr(), N, emu.ch, emu.stream()))
        for r, (qx, sx, qw, sw) in enumerate(ops):
            ext._check(emu.lib.ql_tp_emu_set_rank(r))
            ext._check(emu.lib.ql_tp_gemm_push(qx.data_ptr(), sx.data_ptr(), qw.data_ptr(), sw.data_ptr(), None,
                                               M, N, K, ext.fmt_of(dtype), emu.ch, emu.stream()))
        ext._check(emu.lib.ql_tp_reduce_quant_gather(M, N, ext.fmt_of(dtype), 1, emu.ch, emu.stream()))

    one_round()                                            # eager
    torch.cuda.synchronize()
    assert all(torch.equal(out[r], y) for r in range(W))
    st = torch.cuda.Stream()
    st.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(st):
        gr = torch.cuda.CUDAGraph()
        with torch.cuda.graph(gr, stream=st):
            one_round()
    torch.cuda.current_stream().wait_stream(st)
    for _ in range(3):
        out.zero_()
        for r in range(W):
            emu.buf(r, 1, M * N).fill_(0)
        gr.replay()
        torch.cuda.synchronize()
        assert all(torch.equal(out[r], y) for r in range(W))
        for r in range(W):
            assert torch.equal(emu.buf(r, 1, M * N), q.view(torch.uint8).view(-1))
            assert torch.equal(emu.buf(r, 2, M * 4).view(torch.float32), s)


def test_capacity_and_argument_checks(ext, emu):
    lib = emu.lib
    assert lib.ql_tp_max_tokens() == TOKENS and lib.ql_tp_max_hidden() == HIDDEN
    assert lib.ql_tp_reduce_quant_gather(TOKENS + 1, 1024, 0, 0, 0, None) == 5      # QL_ERR_WORKSPACE
    assert lib.ql_tp_reduce_quant_gather(512, HIDDEN + 8, 0, 0, 0, None) == 5
    assert lib.ql_tp_reduce_quant_gather(512, 1020, 0, 0, 0, None) == 2             # N % 8
    assert lib.ql_tp_reduce_quant_gather(512, 1024, 9, 0, 0, None) == 1             # bad format
    assert lib.ql_tp_reduce_quant_gather(0, 1024, 0, 0, 0, None) == 0               # empty: nothing to do
    assert lib.ql_tp_emu_set_rank(emu.world) == 2
    assert lib.ql_tp_reduce_quant_gather(512, 1024, 0, 0, 2, None) == 2   # no such channel, whatever was allocated
    assert lib.ql_tp_channels() == (2 if emu.ch else 1)                # channel 1 exists only with the opt-in at init
    # channel 1 without the opt-in: refused with the hazard's own status, before any launch
    saved = os.environ.get("QL_TP_ALLOW_CONCURRENT")
    os.environ["QL_TP_ALLOW_CONCURRENT"] = "0"
    try:
        before = lib.ql_launch_count()
        assert lib.ql_tp_reduce_quant_gather(512, 1024, 0, 0, 1, None) == 9
        assert lib.ql_tp_allreduce_bf16(512, 1024, 0, ctypes.c_void_p(16), 1024, 1, None) == 9
        assert lib.ql_tp_set_concurrent(1) == 9
        assert lib.ql_launch_count() == before
    finally:
        os.environ["QL_TP_ALLOW_CONCURRENT"] = saved if saved is not None else "0"
    assert lib.ql_ar_allreduce_bf16(ctypes.c_void_p(16), ctypes.c_void_p(16), 8, None) == 5   # one-shot path is not emulated
