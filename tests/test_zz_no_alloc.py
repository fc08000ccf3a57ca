"""ql_linear_static: the quantized linear for callers that cannot allocate a workspace.

Some users of the C ABI have nothing to allocate ql_linear's scratch with: a client bound from
plain C / ctypes has no torch caching allocator, and once the serving engine has handed the rest
of HBM to its KV cache a raw cudaMalloc fails (or stalls the device mid-forward).  For them
ql_linear_static keeps the quantised activations in an 8 MiB array inside the library image.
These tests call it through the raw C ABI (ctypes, device pointers only -- the scratch is never
passed) and require the result of the existing entry point, ql_linear, on the same arguments:
bit-equal for INT8, within the FP8 tolerance of tests/test_gpu_parity.py otherwise.  Every case
makes two calls in a row on one stream with different shapes and data, because the scratch is
shared and still holds the previous call's bytes.

The file name sorts last on purpose: `pytest -x` stops at the first failure, and a failure here
must not hide the parity tests.  It passes alone in a fresh process and after the other files.
"""
import ctypes

import numpy as np
import pytest
import torch

from oracle import oracle as orc

FP_TOL_ULP = 1.0      # of max|y|, as tests/test_gpu_parity.py
DT = {"i8": torch.int8, "f8": torch.float8_e4m3fn}
MIB = 1 << 20


@pytest.fixture(scope="module")
def ext():
    from nanovllm import _ext
    _ext.load()
    return _ext


def test_static_scratch_limit_is_the_documented_one(ext):
    lib = ext.load()
    assert lib.ql_static_scratch_bytes() == 8 * MIB
    # the examples given in include/qlinear_b200.h fit, the next size up does not
    assert lib.ql_workspace_bytes(4096, 1024) <= 8 * MIB
    assert lib.ql_workspace_bytes(2048, 3072) <= 8 * MIB
    assert lib.ql_workspace_bytes(4096, 3072) > 8 * MIB
    assert "scratch" in lib.ql_status_string(8).decode()


def test_static_bad_arguments_keep_their_own_codes(ext):
    """Validation is ql_linear's, also when the problem is over the limit (no GPU needed)."""
    lib = ext.load()
    fake = ctypes.c_void_p(0x1000)
    # K % 128 != 0 (reference quantization.py:328), small and oversized
    assert lib.ql_linear_static(fake, 100, fake, fake, None, fake, 8, 4, 8, 100, 0, 0, None) == 2
    assert lib.ql_linear_static(fake, 1 << 20, fake, fake, None, fake, 8, 4096, 8, (1 << 20) - 28, 0, 0, None) == 2
    # bad format on an oversized problem
    assert lib.ql_linear_static(fake, 4096, fake, fake, None, fake, 8, 8192, 8, 4096, 7, 0, None) == 1
    # NULL input on an oversized problem
    assert lib.ql_linear_static(None, 4096, fake, fake, None, fake, 8, 8192, 8, 4096, 0, 0, None) == 7
    # empty problems succeed without touching the device
    assert lib.ql_linear_static(fake, 128, fake, fake, None, fake, 8, 0, 8, 128, 0, 0, None) == 0


def _problem(ext, M, N, K, kind, group, seed, bias):
    g = torch.Generator().manual_seed(seed)
    x = torch.randn(M, K, generator=g).to(torch.bfloat16).cuda()
    w = (torch.randn(N, K, generator=g) * 0.02).to(torch.bfloat16).cuda()
    b = torch.randn(N, generator=g).to(torch.bfloat16).cuda() if bias else None
    if group:
        qw, sw = ext.quant_groups(w, group, DT[kind])
    else:
        qw, sw = ext.quant_rows(w, DT[kind])
    return x, qw, sw, b


def _call_static(ext, x, qw, sw, b, kind, group):
    """The raw C-ABI call a torch-less client would make: pointers, sizes, the stream."""
    M, K = x.shape
    N = qw.shape[0]
    y = torch.empty(M, N, dtype=torch.bfloat16, device=x.device)
    stream = torch.cuda.current_stream().cuda_stream
    st = ext.load().ql_linear_static(x.data_ptr(), x.stride(0), qw.data_ptr(), sw.data_ptr(),
                                     None if b is None else b.data_ptr(), y.data_ptr(), N, M, N, K,
                                     ext.fmt_of(DT[kind]), group, stream)
    assert st == 0, ext.load().ql_status_string(st).decode()
    return y


def _assert_same(y, ref, kind, what):
    a, r = orc.t2n(y), orc.t2n(ref)
    if kind == "i8":
        assert np.array_equal(a, r), f"{what}: INT8 output differs from ql_linear"
        return
    af, rf = orc.bf16_decode(a).astype(np.float64), orc.bf16_decode(r).astype(np.float64)
    amax = np.abs(rf).max()
    ulp_at_max = 2.0 ** (np.floor(np.log2(amax)) - 7)
    assert np.abs(af - rf).max() <= FP_TOL_ULP * ulp_at_max, what


# (first call, second call): the second is issued right behind the first, no synchronisation,
# into a scratch that still holds the first call's activations
CASES = [
    # prefill tiles, then a smaller ragged problem that leaves stale bytes behind its own
    ("i8", 0, (1024, 1536, 1024, True), (300, 512, 384, False)),
    # the largest 0.6B shape that fits (4096 tokens x K 1024), then a decode step
    ("i8", 0, (4096, 1024, 1024, False), (8, 4096, 1024, False)),
    # decode first (two-launch path through the scratch for a long K), then prefill
    ("i8", 0, (16, 1024, 4096, False), (512, 1024, 2048, True)),
    ("f8", 0, (768, 1024, 1024, True), (200, 768, 512, False)),
    ("f8", 128, (512, 1024, 1024, False), (64, 512, 256, True)),
    ("i8", 128, (640, 768, 512, False), (33, 256, 384, False)),
]


@pytest.mark.gpu
@pytest.mark.parametrize("kind,group,first,second", CASES)
def test_static_matches_ql_linear_on_two_calls_in_a_row(ext, kind, group, first, second):
    assert torch.cuda.is_available(), "gpu tests need a CUDA device"
    lib = ext.load()
    assert lib.ql_device_check() == 0, "device is not sm_100"
    probs = [_problem(ext, M, N, K, kind, group, 100 + i, bias)
             for i, (M, N, K, bias) in enumerate((first, second))]
    before = lib.ql_launch_count()
    outs = [_call_static(ext, *p, kind, group) for p in probs]          # back to back
    assert lib.ql_launch_count() > before, "no kernel of the library ran"
    torch.cuda.synchronize()
    refs = [ext.linear(x, qw, sw, b, DT[kind], group) for (x, qw, sw, b) in probs]
    torch.cuda.synchronize()
    for i, (y, ref) in enumerate(zip(outs, refs)):
        _assert_same(y, ref, kind, f"{kind} g{group} call {i + 1} {(first, second)[i]}")
    # and once more in the opposite order: the results do not depend on what the scratch held
    outs2 = [_call_static(ext, *p, kind, group) for p in reversed(probs)]
    torch.cuda.synchronize()
    for y, ref in zip(outs2, reversed(refs)):
        _assert_same(y, ref, kind, f"{kind} g{group} reversed order")


@pytest.mark.gpu
def test_static_refuses_what_does_not_fit(ext):
    assert torch.cuda.is_available(), "gpu tests need a CUDA device"
    lib = ext.load()
    M, N, K = 4096, 256, 3072                      # 12 MiB of codes: over the 8 MiB limit
    x, qw, sw, _ = _problem(ext, M, N, K, "i8", 0, 7, False)
    y = torch.full((M, N), 3.0, dtype=torch.bfloat16, device="cuda")
    before = lib.ql_launch_count()
    st = lib.ql_linear_static(x.data_ptr(), K, qw.data_ptr(), sw.data_ptr(), None, y.data_ptr(), N,
                              M, N, K, ext.QL_INT8, 0, torch.cuda.current_stream().cuda_stream)
    assert st == 8 and "scratch" in lib.ql_status_string(st).decode()
    assert lib.ql_launch_count() == before, "a refused call must not launch anything"
    torch.cuda.synchronize()
    assert bool((y == 3.0).all()), "a refused call must not write the output"
    # the same tokens in two halves fit and equal the single ql_linear call
    ref = ext.linear(x, qw, sw, None, torch.int8, 0)
    halves = [_call_static(ext, x[i:i + 2048], qw, sw, None, "i8", 0) for i in (0, 2048)]
    torch.cuda.synchronize()
    assert torch.equal(torch.cat(halves), ref)
