"""Decode-regime chain of quantized linears in one persistent launch (ql_chain_*,
csrc/decode_chain.cu) -- the M <= 16 form of the reference's per_row_quant + per_row_matmul
pair (nanovllm/utils/quantization.py:310-346) applied linear after linear (qwen3.py:80,90,116,118).

Bars:
  * INT8: every op's output BIT-EXACT vs the CPU oracle (quant_rows -> gemm_rows) applied to the
    chain's own input of that op, and vs ql_linear on the same input;
  * FP8: every op within 1 bf16 ulp of max|y| of the oracle (fp32 summation order differs);
  * the kernel re-arms its own counters: a second run and a CUDA-graph replay give the same bits.
"""
import ctypes

import numpy as np
import pytest
import torch

from oracle import oracle as orc

DT = {"i8": torch.int8, "f8": torch.float8_e4m3fn}
OF = {"i8": orc.INT8, "f8": orc.FP8}

# (N, K) per op; op i > 0 consumes the first K columns of op i-1's output unless marked "own"
QWEN06 = [(4096, 1024), (1024, 2048), (6144, 1024), (1024, 3072)]
QWEN8 = [(6144, 4096), (4096, 4096), (24576, 4096), (4096, 12288)]
ODD = [(264, 384), (128, 256), (1032, 128), (8, 1024)]


@pytest.fixture(params=["tcgen05", "mma_sync"], autouse=True)
def chain_kernel(request, monkeypatch):
    """Both forms of the chain: decode_chain_tc_kernel (every K % 512 == 0; the default) and
    decode_chain_kernel (mma.sync; also the fallback for other K).  QL_CHAIN_TC is read by
    ql_chain_prepare."""
    monkeypatch.setenv("QL_CHAIN_TC", "1" if request.param == "tcgen05" else "0")
    return request.param


def make_ops(shapes, M, fmt, seed, own=(), bias_at=(), x_stride_pad=0):
    from nanovllm import _ext
    g = torch.Generator(device="cuda").manual_seed(seed)
    ops, xs = [], []
    for i, (N, K) in enumerate(shapes):
        w = (torch.randn(N, K, generator=g, device="cuda") * 0.05).to(torch.bfloat16)
        qw, sw = _ext.quant_rows(w, DT[fmt])
        b = (torch.randn(N, generator=g, device="cuda") * 0.1).to(torch.bfloat16) if i in bias_at else None
        x = None
        if i == 0 or i in own:
            buf = torch.randn(M, K + x_stride_pad, generator=g, device="cuda").to(torch.bfloat16)
            x = buf[:, :K]
        ops.append((qw, sw, b, x))
        xs.append(x)
    return ops, xs


def check_chain(chain, ops, xs, M, fmt):
    from nanovllm import _ext
    outs = [o.clone() for o in chain.outputs]
    for i, (qw, sw, b, _) in enumerate(ops):
        N, K = qw.shape
        x = xs[i] if xs[i] is not None else outs[i - 1][:, :K]
        qx, sx = orc.quant_rows(orc.t2n(x.contiguous()), OF[fmt])
        yo = orc.gemm_rows(qx, sx, orc.t2n(qw), orc.t2n(sw), None if b is None else orc.t2n(b), OF[fmt])
        got = orc.t2n(outs[i])
        if fmt == "i8":
            assert np.array_equal(got, yo), f"op {i} ({N}x{K}) differs from the oracle"
            y_lin = _ext.linear(x, qw, sw, b, DT[fmt])
            assert torch.equal(outs[i], y_lin), f"op {i} differs from ql_linear"
        else:
            ref = orc.bf16_decode(yo)
            err = np.abs(orc.bf16_decode(got) - ref).max()
            tol = 2.0 ** (np.floor(np.log2(max(np.abs(ref).max(), 1e-30))) - 7)
            assert err <= tol, f"op {i} ({N}x{K}): err {err} > {tol}"
    return outs


@pytest.mark.gpu
@pytest.mark.parametrize("fmt", ["i8", "f8"])
@pytest.mark.parametrize("M", [1, 5, 16])
def test_chain_qwen06_two_layers(fmt, M):
    from nanovllm import _ext
    shapes = QWEN06 + QWEN06
    ops, xs = make_ops(shapes, M, fmt, seed=M)
    chain = _ext.DecodeChain(ops, M, DT[fmt])
    n0 = _ext.launch_count()
    chain.run()
    assert _ext.launch_count() - n0 == 1                    # ONE launch for 8 linears
    torch.cuda.synchronize()
    first = check_chain(chain, ops, xs, M, fmt)
    chain.run()                                             # counters re-armed by the kernel itself
    torch.cuda.synchronize()
    for a, b in zip(first, chain.outputs):
        assert torch.equal(a, b)


@pytest.mark.gpu
@pytest.mark.parametrize("fmt", ["i8", "f8"])
def test_chain_qwen8_layer_k_slices(fmt):
    """K = 12288: several K slices per row block meet in L2 (last arriver finishes)."""
    from nanovllm import _ext
    M = 16
    ops, xs = make_ops(QWEN8, M, fmt, seed=7, bias_at=(1, 3))
    chain = _ext.DecodeChain(ops, M, DT[fmt])
    chain.run()
    torch.cuda.synchronize()
    check_chain(chain, ops, xs, M, fmt)


@pytest.mark.gpu
@pytest.mark.parametrize("M", [3, 16])
def test_chain_odd_shapes_own_inputs_bias_stride(M):
    """Ragged sizes (N % 8 == 0 only, K = 128 .. 1024), an op in the middle with its own strided
    input (amax scanned in-kernel instead of taken from the previous epilogue), biases."""
    from nanovllm import _ext
    ops, xs = make_ops(ODD, M, "i8", seed=11 + M, own=(2,), bias_at=(0, 2), x_stride_pad=24)
    chain = _ext.DecodeChain(ops, M, torch.int8)
    chain.run()
    torch.cuda.synchronize()
    check_chain(chain, ops, xs, M, "i8")


@pytest.mark.gpu
def test_chain_graph_replay_and_new_input():
    from nanovllm import _ext
    M = 8
    ops, xs = make_ops(QWEN06, M, "i8", seed=3)
    chain = _ext.DecodeChain(ops, M, torch.int8)
    s = torch.cuda.Stream()
    s.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(s):
        chain.run()
        gr = torch.cuda.CUDAGraph()
        with torch.cuda.graph(gr, stream=s):
            chain.run()
    torch.cuda.current_stream().wait_stream(s)
    for rep in range(3):
        xs[0].copy_(torch.randn(M, xs[0].shape[1], device="cuda").to(torch.bfloat16))
        gr.replay()
        torch.cuda.synchronize()
        check_chain(chain, ops, xs, M, "i8")


@pytest.mark.gpu
def test_chain_rejects_bad_arguments():
    from nanovllm import _ext
    lib = _ext.load()
    ops, _ = make_ops(QWEN06[:1], 4, "i8", seed=1)
    qw, sw, _, x = ops[0]
    arr = (_ext._ChainOp * 1)(_ext._ChainOp(qw.data_ptr(), sw.data_ptr(), None, x.data_ptr(), x.data_ptr(),
                                            1024, 4096, 4096, 1024))
    assert lib.ql_chain_workspace_bytes(ctypes.addressof(arr), 1, 17) == 0        # M > 16
    assert lib.ql_chain_workspace_bytes(ctypes.addressof(arr), 0, 4) == 0
    nbytes = lib.ql_chain_workspace_bytes(ctypes.addressof(arr), 1, 4)
    assert nbytes > 0
    ws = torch.empty(nbytes, dtype=torch.uint8, device="cuda")
    plan = ctypes.create_string_buffer(128)
    st = torch.cuda.current_stream().cuda_stream
    assert lib.ql_chain_prepare(ctypes.addressof(arr), 1, 4, 0, 128, ws.data_ptr(), nbytes,
                                ctypes.addressof(plan), st) != 0                   # per-group: not here
    assert lib.ql_chain_prepare(ctypes.addressof(arr), 1, 4, 7, 0, ws.data_ptr(), nbytes,
                                ctypes.addressof(plan), st) != 0                   # bad format
    assert lib.ql_chain_prepare(ctypes.addressof(arr), 1, 4, 0, 0, ws.data_ptr(), nbytes - 256,
                                ctypes.addressof(plan), st) != 0                   # short workspace
    assert lib.ql_chain_run(ctypes.addressof(plan), ws.data_ptr(), st) != 0        # never prepared
