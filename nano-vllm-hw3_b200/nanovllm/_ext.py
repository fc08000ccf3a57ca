"""ctypes binding of libqlinear_b200.so (C ABI: include/qlinear_b200.h) + the torch glue.

This is the "thin torch extension" of the quantized-linear path: torch owns every tensor and
the stream, the library gets raw device pointers, sizes and ``cudaStream_t``.  Argument
checks raise ``AssertionError`` where the reference asserts (nanovllm/utils/quantization.py
:151-152, :203-205, :328) and ``RuntimeError`` for everything the library itself reports.

There is NO fallback: if the shared library is missing or the device is not a B200
(compute capability 10.x) the first call raises.
"""
from __future__ import annotations

import ctypes
import os

import torch

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "_lib", "libqlinear_b200.so")

QL_INT8, QL_FP8_E4M3 = 0, 1
QL_ABI_VERSION = 9
_FMT = {torch.int8: QL_INT8, torch.float8_e4m3fn: QL_FP8_E4M3}

_c_i64, _c_int, _c_vp, _c_sz = ctypes.c_int64, ctypes.c_int, ctypes.c_void_p, ctypes.c_size_t

# name -> (restype, argtypes); mirrors include/qlinear_b200.h one to one
SIGNATURES = {
    "ql_version": (_c_int, []),
    "ql_status_string": (ctypes.c_char_p, [_c_int]),
    "ql_device_check": (_c_int, []),
    "ql_launch_count": (ctypes.c_uint64, []),
    "ql_debug_set_trace": (None, [_c_vp]),
    "ql_quant_rows": (_c_int, [_c_vp, _c_i64, _c_i64, _c_i64, _c_int, _c_vp, _c_vp, _c_vp]),
    "ql_quant_groups": (_c_int, [_c_vp, _c_i64, _c_i64, _c_i64, _c_int, _c_vp, _c_vp, _c_vp]),
    "ql_quant_tensor": (_c_int, [_c_vp, _c_i64, _c_int, _c_vp, _c_vp, _c_vp, _c_sz, _c_vp]),
    "ql_gemm_w8a8": (_c_int, [_c_vp, _c_vp, _c_vp, _c_vp, _c_vp, _c_vp, _c_i64, _c_i64, _c_i64,
                              _c_i64, _c_int, _c_i64, _c_vp]),
    "ql_gemm_i8_acc": (_c_int, [_c_vp, _c_vp, _c_vp, _c_i64, _c_i64, _c_i64, _c_vp]),
    "ql_gemv_w8a8": (_c_int, [_c_vp, _c_i64, _c_vp, _c_vp, _c_vp, _c_vp, _c_i64, _c_i64, _c_i64,
                              _c_i64, _c_int, _c_i64, _c_vp]),
    "ql_linear": (_c_int, [_c_vp, _c_i64, _c_vp, _c_vp, _c_vp, _c_vp, _c_i64, _c_i64, _c_i64,
                           _c_i64, _c_int, _c_i64, _c_vp, _c_sz, _c_vp]),
    "ql_rmsnorm_quant": (_c_int, [_c_vp, _c_i64, _c_vp, _c_vp, _c_vp, ctypes.c_float, _c_i64, _c_i64,
                                  _c_int, _c_vp, _c_vp, _c_vp, _c_vp]),
    "ql_silu_mul_quant": (_c_int, [_c_vp, _c_i64, _c_i64, _c_i64, _c_int, _c_vp, _c_vp, _c_vp, _c_vp]),
    "ql_ar_handle_bytes": (_c_sz, []),
    "ql_ar_can_access_peers": (_c_int, [_c_int]),
    "ql_ar_init": (_c_int, [_c_sz, _c_sz, _c_sz, _c_int, _c_int, _c_vp]),
    "ql_ar_open": (_c_int, [_c_vp]),
    "ql_ar_destroy": (_c_int, []),
    "ql_ar_max_bytes": (_c_sz, []),
    "ql_ar_allreduce_bf16": (_c_int, [_c_vp, _c_vp, _c_i64, _c_vp]),
    "ql_ar_rmsnorm_quant": (_c_int, [_c_vp, _c_vp, _c_vp, _c_vp, ctypes.c_float, _c_i64, _c_i64, _c_int,
                                     _c_vp, _c_vp, _c_vp]),
    "ql_ar_init_local": (_c_int, [_c_sz, _c_sz, _c_sz, _c_int]),
    "ql_tp_emu_set_rank": (_c_int, [_c_int]),
    "ql_tp_max_tokens": (_c_sz, []),
    "ql_tp_max_hidden": (_c_sz, []),
    "ql_tp_channels": (_c_int, []),
    "ql_tp_set_concurrent": (_c_int, [_c_int]),
    "ql_tp_buffer": (_c_vp, [_c_int, _c_int]),
    "ql_tp_gemm_push": (_c_int, [_c_vp, _c_vp, _c_vp, _c_vp, _c_vp, _c_i64, _c_i64, _c_i64, _c_int, _c_int, _c_vp]),
    "ql_tp_reduce_quant_gather": (_c_int, [_c_i64, _c_i64, _c_int, _c_int, _c_int, _c_vp]),
    "ql_tp_allreduce_bf16": (_c_int, [_c_i64, _c_i64, _c_int, _c_vp, _c_i64, _c_int, _c_vp]),
    "ql_workspace_bytes": (_c_sz, [_c_i64, _c_i64]),
    "ql_static_scratch_bytes": (_c_sz, []),
    "ql_linear_static": (_c_int, [_c_vp, _c_i64, _c_vp, _c_vp, _c_vp, _c_vp, _c_i64, _c_i64, _c_i64,
                                  _c_i64, _c_int, _c_i64, _c_vp]),
    "ql_chain_workspace_bytes": (_c_sz, [_c_vp, _c_int, _c_i64]),
    "ql_chain_prepare": (_c_int, [_c_vp, _c_int, _c_i64, _c_int, _c_i64, _c_vp, _c_sz, _c_vp, _c_vp]),
    "ql_chain_run": (_c_int, [_c_vp, _c_vp, _c_vp]),
    "ql_linear_path": (ctypes.c_char_p, [_c_i64, _c_i64, _c_i64, _c_i64]),
}

_lib = None


def load():
    """dlopen the library (once).  Raises RuntimeError when it has not been built."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise RuntimeError(
                f"{LIB_PATH} is missing: build it with `python nano-vllm-hw3_b200/build.py` "
                "(nvcc, sm_100a).  The quantized-linear path has no CPU/Triton fallback.")
        lib = ctypes.CDLL(LIB_PATH)
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(lib, name)
            fn.restype, fn.argtypes = res, args
        if lib.ql_version() != QL_ABI_VERSION:
            raise RuntimeError(f"libqlinear_b200.so ABI {lib.ql_version()} != binding {QL_ABI_VERSION}")
        _lib = lib
    return _lib


def _check(status: int):
    if status != 0:
        raise RuntimeError(f"qlinear_b200: {load().ql_status_string(status).decode()} (status {status})")


def fmt_of(dtype: torch.dtype) -> int:
    assert dtype in [torch.int8, torch.float8_e4m3fn]      # reference quantization.py:151,203,353
    return _FMT[dtype]


def _stream(t: torch.Tensor) -> int:
    return torch._C._cuda_getCurrentRawStream(t.device.index)


class _on_device:
    """``with torch.cuda.device(t.device)`` without its cost when the device is already current
    (the engine runs one process per GPU, model_runner.py:29)."""

    __slots__ = ("index", "prev")

    def __init__(self, t: torch.Tensor):
        self.index = t.device.index

    def __enter__(self):
        cur = torch.cuda.current_device()
        self.prev = cur if cur != self.index else -1
        if self.prev >= 0:
            torch.cuda.set_device(self.index)

    def __exit__(self, *exc):
        if self.prev >= 0:
            torch.cuda.set_device(self.prev)
        return False


def _ptr(t):
    return None if t is None else t.data_ptr()


def _cuda_rowmajor(t: torch.Tensor, what: str):
    if not t.is_cuda:
        raise RuntimeError(f"qlinear_b200: {what} must be a CUDA tensor (no CPU path exists)")
    assert t.stride(-1) == 1, f"{what} must be contiguous along its last dimension"


# ------------------------------------------------------------------------------ entry points

@torch.no_grad()
def quant_rows(x: torch.Tensor, dtype: torch.dtype):
    """ql_quant_rows: x bf16 [M,K] -> (q dtype [M,K], s fp32 [M])."""
    fmt = fmt_of(dtype)
    assert x.dim() == 2                                      # reference :152
    assert x.dtype == torch.bfloat16, "activations/weights are bf16 (reference default dtype)"
    _cuda_rowmajor(x, "x")
    M, K = x.shape
    q = torch.empty((M, K), dtype=dtype, device=x.device)
    s = torch.empty((M,), dtype=torch.float32, device=x.device)
    with _on_device(x):
        _check(load().ql_quant_rows(_ptr(x), M, K, x.stride(0) if M > 1 else K, fmt, _ptr(q),
                                    _ptr(s), _stream(x)))
    return q, s


@torch.no_grad()
def quant_groups(w: torch.Tensor, group_size: int, dtype: torch.dtype):
    """ql_quant_groups: w bf16 [N,K] -> (q dtype [N,K], s fp32 [N, K/group_size])."""
    fmt = fmt_of(dtype)
    assert w.dim() == 2                                      # reference :204
    assert w.size(1) % group_size == 0, \
        f"K dim ({w.size(1)}) must be divisible by group_size ({group_size})"   # reference :205
    assert w.dtype == torch.bfloat16
    w = w.contiguous()
    _cuda_rowmajor(w, "w")
    N, K = w.shape
    q = torch.empty((N, K), dtype=dtype, device=w.device)
    s = torch.empty((N, K // group_size), dtype=torch.float32, device=w.device)
    with _on_device(w):
        _check(load().ql_quant_groups(_ptr(w), N, K, group_size, fmt, _ptr(q), _ptr(s), _stream(w)))
    return q, s


@torch.no_grad()
def quant_tensor(x: torch.Tensor, dtype: torch.dtype):
    """ql_quant_tensor: x bf16 (any shape) -> (q dtype like x, scale 0-dim bf16), reference :166-191."""
    fmt = fmt_of(dtype)
    assert x.dtype == torch.bfloat16
    x = x.contiguous()
    _cuda_rowmajor(x, "x")
    q = torch.empty(x.shape, dtype=dtype, device=x.device)
    scale = torch.empty((), dtype=torch.bfloat16, device=x.device)
    ws = torch.empty((1024,), dtype=torch.float32, device=x.device)
    with _on_device(x):
        _check(load().ql_quant_tensor(_ptr(x), x.numel(), fmt, _ptr(q), _ptr(scale), _ptr(ws),
                                      ws.numel() * 4, _stream(x)))
    return q, scale


def _prep_linear(x, qw, sw, bias, group_size):
    assert x.dtype == torch.bfloat16
    _cuda_rowmajor(x, "x")
    K = x.shape[-1]
    N = qw.shape[0]
    assert qw.dim() == 2 and qw.shape[1] == K and qw.is_contiguous()
    assert K % 128 == 0, f"K({K}) must be multiple of BLK_K(128)"          # reference :328
    assert sw.dtype == torch.float32 and sw.is_contiguous()
    if group_size:
        assert K % group_size == 0
        assert tuple(sw.shape) == (N, K // group_size)
    else:
        assert sw.numel() == N
    if bias is not None:
        assert bias.dtype == torch.bfloat16 and bias.numel() == N and bias.is_contiguous()
    if x.dim() == 2:                                         # the engine's token-major activations
        x2, y_shape = x, None
    else:
        y_shape = list(x.shape[:-1]) + [N]
        x2 = x.reshape(-1, K)                                # reference :322 (view when possible)
    if x2.stride(-1) != 1 or (x2.shape[0] > 1 and x2.stride(0) % 8 != 0) or x2.data_ptr() % 16:
        x2 = x2.contiguous()
    return x2, y_shape, N, K


# One scratch buffer per (device, stream) instead of a torch.empty per call: every use is ordered
# on that stream (the next call's quantiser starts behind griddepcontrol.wait, i.e. after the GEMM
# that read the buffer has finished; tests/test_gpu_parity.py::test_linear_shared_workspace_back_to_back
# pins that).  A buffer is replaced only when a larger one is needed, and the old one is released
# after its stream has drained, so a pointer handed to a kernel never dangles and nothing piles up.
# During CUDA-graph capture the scratch comes from the graph's own pool instead (graphs captured on
# one stream may be replayed concurrently).
_WS = {}
_get_raw_stream = getattr(torch._C, "_cuda_getCurrentRawStream", None)     # private but cheap; public fallback below


def _raw_stream(index: int) -> int:
    if _get_raw_stream is not None:
        return _get_raw_stream(index)
    return torch.cuda.current_stream(index).cuda_stream


def _workspace(dev, index: int, stream: int, nbytes: int):
    if torch.cuda.is_current_stream_capturing():
        return torch.empty((nbytes,), dtype=torch.uint8, device=dev)
    key = (index, stream)
    ws = _WS.get(key)
    if ws is None or ws.numel() < nbytes:
        if ws is not None:
            # the kernels still reading the outgrown buffer are all on this stream: drain it, then
            # let the caching allocator have the block back (growth is rare: sizes only go up)
            torch.cuda.current_stream(index).synchronize()
            del _WS[key]
            ws = None
        ws = torch.empty((max(nbytes, 1 << 20),), dtype=torch.uint8, device=dev)
        _WS[key] = ws
    return ws


def workspace_bytes(M: int, K: int) -> int:
    """ql_workspace_bytes, mirrored (qx [M,K] rounded to 256 B + sx [M] rounded to 256 B)."""
    return ((M * K + 255) // 256 + (M * 4 + 255) // 256) * 256 if M > 0 and K > 0 else 0


@torch.no_grad()
def linear(x, qw, sw, bias, dtype, group_size: int = 0, out: torch.Tensor | None = None):
    """ql_linear: one whole quantized linear, y = bf16[..., N].  group_size 0 = per-row scales."""
    fmt = fmt_of(dtype)
    assert qw.dtype == dtype
    x2, y_shape, N, K = _prep_linear(x, qw, sw, bias, group_size)
    M = x2.shape[0]
    dev = x2.device
    index = dev.index
    y = out if out is not None else torch.empty((M, N), dtype=torch.bfloat16, device=dev)
    lib = load()
    switch = index != torch.cuda.current_device()
    if switch:
        prev = torch.cuda.current_device()
        torch.cuda.set_device(index)
    try:
        stream = _raw_stream(index)
        nbytes = workspace_bytes(M, K)
        ws = _workspace(dev, index, stream, nbytes) if nbytes else None
        _check(lib.ql_linear(x2.data_ptr(), x2.stride(0) if M > 1 else K, qw.data_ptr(), sw.data_ptr(),
                             None if bias is None else bias.data_ptr(), y.data_ptr(),
                             y.stride(0) if M > 1 else N, M, N, K, fmt, group_size,
                             None if ws is None else ws.data_ptr(), nbytes, stream))
    finally:
        if switch:
            torch.cuda.set_device(prev)
    if out is not None:
        return out
    return y if y_shape is None else y.view(y_shape)


@torch.no_grad()
def gemm_w8a8(qx, sx, qw, sw, bias, dtype, group_size: int = 0, out: torch.Tensor | None = None):
    """ql_gemm_w8a8 on pre-quantised activations (tcgen05 path regardless of M)."""
    fmt = fmt_of(dtype)
    assert qx.dtype == dtype and qw.dtype == dtype and qx.is_contiguous() and qw.is_contiguous()
    M, K = qx.shape
    N = qw.shape[0]
    assert K % 128 == 0, f"K({K}) must be multiple of BLK_K(128)"
    y = out if out is not None else torch.empty((M, N), dtype=torch.bfloat16, device=qx.device)
    with _on_device(qx):
        _check(load().ql_gemm_w8a8(_ptr(qx), _ptr(sx), _ptr(qw), _ptr(sw), _ptr(bias), _ptr(y), N,
                                   M, N, K, fmt, group_size, _stream(qx)))
    return y


@torch.no_grad()
def gemm_i8_acc(qx, qw):
    """ql_gemm_i8_acc: raw int32 accumulators of the tcgen05 INT8 kernel (parity entry point)."""
    assert qx.dtype == torch.int8 and qw.dtype == torch.int8
    assert qx.is_contiguous() and qw.is_contiguous()
    M, K = qx.shape
    N = qw.shape[0]
    assert K % 128 == 0
    acc = torch.empty((M, N), dtype=torch.int32, device=qx.device)
    with _on_device(qx):
        _check(load().ql_gemm_i8_acc(_ptr(qx), _ptr(qw), _ptr(acc), M, N, K, _stream(qx)))
    return acc


@torch.no_grad()
def gemv_w8a8(x, qw, sw, bias, dtype, group_size: int = 0):
    """ql_gemv_w8a8: fused quantise + GEMV, M <= 16."""
    fmt = fmt_of(dtype)
    x2, y_shape, N, K = _prep_linear(x, qw, sw, bias, group_size)
    M = x2.shape[0]
    y = torch.empty((M, N), dtype=torch.bfloat16, device=x.device)
    with _on_device(x):
        _check(load().ql_gemv_w8a8(_ptr(x2), x2.stride(0) if M > 1 else K, _ptr(qw), _ptr(sw),
                                   _ptr(bias), _ptr(y), N, M, N, K, fmt, group_size, _stream(x)))
    return y if y_shape is None else y.view(y_shape)


@torch.no_grad()
def rmsnorm_quant(x, weight, eps: float, residual=None, dtype=None, want_bf16: bool = False,
                  residual_out=None):
    """ql_rmsnorm_quant.  x bf16 [M,K]; residual bf16 [M,K] or None: x + residual is written to
    ``residual_out`` (bf16 [M,K] dense) or, when that is None, IN PLACE over ``residual`` (the
    engine's residual stream).  Returns (q, s, h): q/s None when dtype is None, h None unless
    want_bf16 (or dtype is None)."""
    assert x.dtype == torch.bfloat16 and x.dim() == 2 and weight.dtype == torch.bfloat16
    _cuda_rowmajor(x, "x")
    M, K = x.shape
    assert weight.numel() == K and weight.is_contiguous()
    if residual is not None:
        assert residual.shape == x.shape and residual.dtype == torch.bfloat16 and residual.is_contiguous()
        if residual_out is None:
            residual_out = residual
        assert residual_out.shape == x.shape and residual_out.dtype == torch.bfloat16 and \
            residual_out.is_contiguous()
    else:
        assert residual_out is None
    fmt = fmt_of(dtype) if dtype is not None else 0
    q = torch.empty((M, K), dtype=dtype, device=x.device) if dtype is not None else None
    s = torch.empty((M,), dtype=torch.float32, device=x.device) if dtype is not None else None
    h = torch.empty((M, K), dtype=torch.bfloat16, device=x.device) if (want_bf16 or dtype is None) else None
    with _on_device(x):
        _check(load().ql_rmsnorm_quant(_ptr(x), x.stride(0) if M > 1 else K, _ptr(residual), _ptr(residual_out),
                                       _ptr(weight), float(eps), M, K, fmt, _ptr(q), _ptr(s), _ptr(h),
                                       _stream(x)))
    return q, s, h


@torch.no_grad()
def silu_mul_quant(gate_up, dtype=None, want_bf16: bool = False):
    """ql_silu_mul_quant.  gate_up bf16 [M, 2I] -> (q [M,I], s [M], h bf16 [M,I] or None)."""
    assert gate_up.dtype == torch.bfloat16 and gate_up.dim() == 2 and gate_up.shape[1] % 2 == 0
    _cuda_rowmajor(gate_up, "gate_up")
    M, I = gate_up.shape[0], gate_up.shape[1] // 2
    fmt = fmt_of(dtype) if dtype is not None else 0
    q = torch.empty((M, I), dtype=dtype, device=gate_up.device) if dtype is not None else None
    s = torch.empty((M,), dtype=torch.float32, device=gate_up.device) if dtype is not None else None
    h = torch.empty((M, I), dtype=torch.bfloat16, device=gate_up.device) if (want_bf16 or dtype is None) else None
    with _on_device(gate_up):
        _check(load().ql_silu_mul_quant(_ptr(gate_up), gate_up.stride(0) if M > 1 else 2 * I, M, I, fmt,
                                        _ptr(q), _ptr(s), _ptr(h), _stream(gate_up)))
    return q, s, h


class _ChainOp(ctypes.Structure):
    """ql_chain_op (include/qlinear_b200.h)."""
    _fields_ = [("qw", _c_vp), ("sw", _c_vp), ("bias", _c_vp), ("x", _c_vp), ("y", _c_vp),
                ("ldx", _c_i64), ("ldy", _c_i64), ("N", _c_i64), ("K", _c_i64)]


class DecodeChain:
    """ql_chain_*: a chain of per-row W8A8 linears at M <= 16 tokens in ONE persistent launch.

    ``ops`` is a list of ``(qw, sw, bias, x)``: ``x`` is a bf16 [M, K] tensor (row stride free) or
    None = "the first K columns of the previous op's output" (how o_proj / down_proj consume the
    leading columns of qkv / gate_up in the bench stack, and every layer the previous layer's
    output).  Outputs are allocated here (bf16 [M, N] each, ``self.outputs``) and stay valid for
    the life of the object; ``run()`` launches on the current stream and is graph-capturable."""

    def __init__(self, ops, M: int, dtype: torch.dtype):
        fmt = fmt_of(dtype)
        assert 0 < M <= 16 and len(ops) > 0
        dev = ops[0][0].device
        self.M, self.dtype, self.device = M, dtype, dev
        self.outputs, self._keep = [], []
        arr = (_ChainOp * len(ops))()
        for i, (qw, sw, bias, x) in enumerate(ops):
            assert qw.dtype == dtype and qw.dim() == 2 and qw.is_contiguous() and qw.is_cuda
            N, K = qw.shape
            assert K % 128 == 0, f"K({K}) must be multiple of BLK_K(128)"      # reference :328
            assert sw.dtype == torch.float32 and sw.numel() == N and sw.is_contiguous()
            if bias is not None:
                assert bias.dtype == torch.bfloat16 and bias.numel() == N and bias.is_contiguous()
            if x is not None:
                assert x.dtype == torch.bfloat16 and x.dim() == 2 and x.shape == (M, K) and x.stride(1) == 1
            else:
                assert i > 0 and self.outputs[-1].shape[1] >= K
            y = torch.empty((M, N), dtype=torch.bfloat16, device=dev)
            self.outputs.append(y)
            self._keep.append((qw, sw, bias, x))
            arr[i] = _ChainOp(qw.data_ptr(), sw.data_ptr(), _ptr(bias), _ptr(x), y.data_ptr(),
                              (x.stride(0) if M > 1 else K) if x is not None else 0, N, N, K)
        lib = load()
        with _on_device(qw):
            nbytes = lib.ql_chain_workspace_bytes(ctypes.addressof(arr), len(ops), M)
            if nbytes == 0:
                raise RuntimeError("qlinear_b200: this chain is not supported (ql_chain_workspace_bytes == 0)")
            self._ws = torch.empty((nbytes,), dtype=torch.uint8, device=dev)
            self._plan = ctypes.create_string_buffer(128)
            _check(lib.ql_chain_prepare(ctypes.addressof(arr), len(ops), M, fmt, 0, self._ws.data_ptr(), nbytes,
                                        ctypes.addressof(self._plan), _stream(qw)))

    def run(self):
        with _on_device(self._ws):
            _check(load().ql_chain_run(ctypes.addressof(self._plan), self._ws.data_ptr(), _stream(self._ws)))
        return self.outputs[-1]


def launch_count() -> int:
    """Kernels launched by the library so far in this process."""
    return int(load().ql_launch_count())


def linear_path(M: int, N: int, K: int, group_size: int = 0) -> str:
    return load().ql_linear_path(M, N, K, group_size).decode()
