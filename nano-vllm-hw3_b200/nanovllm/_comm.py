"""Host side of the NVLink peer-memory collectives behind the row-parallel linears
(csrc/allreduce.cu, csrc/tp_collective.cu, csrc/gemm_2cta.cu; C ABI ``ql_ar_*`` / ``ql_tp_*``).

Two regimes replace ``dist.all_reduce(y)`` of ``RowParallelLinear.forward`` (reference
nanovllm/layers/linear.py:150-152):

* decode-sized outputs (<= ``QL_PEER_AR_BYTES``, default 1 MiB): the one-shot push all-reduce
  (``ql_ar_allreduce_bf16``), optionally fused with add + RMSNorm + quant (``ql_ar_rmsnorm_quant``);
* prefill-sized outputs (<= ``QL_TP_TOKENS`` x ``QL_TP_HIDDEN``, default 65536 x 4096): reduce-scatter
  to the token owner -- pushed from the GEMM epilogue (``ql_tp_gemm_push``) or pulled by the owner --
  then the per-token quantiser on the owner's tokens only and an all-gather of codes + scales
  (``ql_tp_reduce_quant_gather``), or of bf16 rows (``ql_tp_allreduce_bf16``).

``torch.distributed`` is only the plumbing: it carries the CUDA IPC handles once at start-up and
agrees on success.  Set-up is collective and failure-tolerant: every stage's status is MIN-reduced
over the ranks, and if any rank fails (no peer access between a GPU pair, cudaIpcOpenMemHandle
refused, out of memory, different sizes per rank) every rank frees what it mapped and returns
``None`` -- the layers then keep using NCCL.  Anything larger than the buffers stays on NCCL too.

``QL_PEER_AR=0`` disables everything here; ``QL_TP_BIG=0`` only the prefill-sized path;
``QL_TP_PUSH=0`` makes the owners pull the partials instead of the GEMM pushing them.

The prefill-sized path has ``ql_tp_channels()`` independent channels (own buffers, flags, call
counters).  A caller that splits a batch into token micro-batches and runs them on separate CUDA
streams, one channel each (``nanovllm.layers.linear.tp_channel``), keeps one collective per channel
in flight, so the wire time of one micro-batch hides under the GEMMs of the other.  That mode has a
KNOWN HANG HAZARD at TP=8 (include/qlinear_b200.h, ql_tp_channels): every call on a channel > 0 is
refused (status 9 -> RuntimeError) unless ``QL_TP_ALLOW_CONCURRENT=1`` is set.
"""
from __future__ import annotations

import ctypes
import os

import torch
import torch.distributed as dist

from nanovllm import _ext

_state = {"tried": False, "comm": None}


class _DevMem:
    """A window of the symmetric area as something ``torch.as_tensor`` can wrap (zero copy)."""

    def __init__(self, ptr: int, nbytes: int):
        self.__cuda_array_interface__ = {"shape": (nbytes,), "typestr": "|u1", "data": (ptr, False),
                                         "version": 2, "strides": None}


def _all_ok(ok: bool) -> bool:
    t = torch.tensor([1 if ok else 0], dtype=torch.int32, device="cuda")
    dist.all_reduce(t, op=dist.ReduceOp.MIN)
    return bool(t.item())


class PeerAllReduce:

    def __init__(self, lib, world: int, rank: int):
        self.lib, self.world, self.rank = lib, world, rank
        self.bytes = int(lib.ql_ar_max_bytes())
        self.tp_tokens = int(lib.ql_tp_max_tokens())
        self.tp_hidden = int(lib.ql_tp_max_hidden())
        self.push = os.environ.get("QL_TP_PUSH", "1") != "0"
        self.channels = int(lib.ql_tp_channels()) if self.tp_tokens else 0
        self._bufs = {}
        if self.tp_tokens:
            dev = torch.device("cuda", torch.cuda.current_device())
            cap = self.tp_tokens * self.tp_hidden
            for ch in range(self.channels):
                for which, nbytes in ((0, cap * 2), (1, cap), (2, self.tp_tokens * 4)):
                    ptr = lib.ql_tp_buffer(which, ch)
                    self._bufs[ch, which] = torch.as_tensor(_DevMem(ptr, nbytes), device=dev)

    @classmethod
    def create(cls, max_bytes: int, tp_tokens: int, tp_hidden: int) -> "PeerAllReduce | None":
        """Collective constructor.  Returns None on EVERY rank if any rank cannot set it up."""
        lib = _ext.load()
        world, rank = dist.get_world_size(), dist.get_rank()
        # every rank must ask for the same area, or remote offsets would be wrong
        sizes = [None] * world
        dist.all_gather_object(sizes, (max_bytes, tp_tokens, tp_hidden, int(lib.ql_tp_channels())))
        if len(set(sizes)) != 1:
            return None
        if not _all_ok(lib.ql_ar_can_access_peers(world) == 0):
            return None
        hb = lib.ql_ar_handle_bytes()
        mine = ctypes.create_string_buffer(hb)
        ok = lib.ql_ar_init(max_bytes, tp_tokens, tp_hidden, world, rank, mine) == 0
        if not _all_ok(ok):
            lib.ql_ar_destroy()
            return None
        handles = [None] * world
        dist.all_gather_object(handles, mine.raw)
        blob = ctypes.create_string_buffer(b"".join(handles), hb * world)
        ok = lib.ql_ar_open(blob) == 0
        if not _all_ok(ok):
            lib.ql_ar_destroy()
            return None
        torch.cuda.synchronize()
        dist.barrier()                       # every peer has mapped every area before first use
        return cls(lib, world, rank)

    def close(self) -> None:
        """Collective: unmap and free (no rank may have a call in flight)."""
        torch.cuda.synchronize()
        dist.barrier()
        self._bufs = {}
        self.lib.ql_ar_destroy()
        _state["comm"] = None

    # ------------------------------------------------------------------ decode-sized: one shot
    def fits(self, numel: int) -> bool:
        return numel % 8 == 0 and 0 < numel * 2 <= self.bytes

    @torch.no_grad()
    def all_reduce(self, y: torch.Tensor) -> torch.Tensor:
        """Sum of every rank's y (fp32, rank order, one rounding) -> fresh bf16 tensor."""
        assert y.dtype == torch.bfloat16 and y.is_contiguous() and self.fits(y.numel())
        out = torch.empty_like(y)
        _ext._check(self.lib.ql_ar_allreduce_bf16(y.data_ptr(), out.data_ptr(), y.numel(),
                                                  torch.cuda.current_stream().cuda_stream))
        return out

    def fits_fused(self, M: int, K: int) -> bool:
        return K % 8 == 0 and K <= 8192 and 0 < M * K * 2 <= self.bytes

    @torch.no_grad()
    def all_reduce_rmsnorm_quant(self, partial, residual, weight, eps: float, dtype):
        """ql_ar_rmsnorm_quant: (q, s, new_residual) from this rank's un-reduced partial [M, K]."""
        assert partial.dtype == torch.bfloat16 and partial.dim() == 2 and partial.is_contiguous()
        M, K = partial.shape
        assert residual.shape == partial.shape and residual.is_contiguous() and self.fits_fused(M, K)
        q = torch.empty((M, K), dtype=dtype, device=partial.device)
        s = torch.empty((M,), dtype=torch.float32, device=partial.device)
        res = torch.empty_like(residual)
        _ext._check(self.lib.ql_ar_rmsnorm_quant(
            partial.data_ptr(), residual.data_ptr(), res.data_ptr(), weight.data_ptr(), float(eps), M, K,
            _ext.fmt_of(dtype), q.data_ptr(), s.data_ptr(), torch.cuda.current_stream().cuda_stream))
        return q, s, res

    # ------------------------------------------------------------------ prefill-sized: reduce-scatter
    def fits_big(self, M: int, N: int) -> bool:
        # token owners hold whole 256-token tiles; the buffers carry that rounding (allreduce.cu)
        return bool(self.tp_tokens) and 0 < M <= self.tp_tokens and N % 8 == 0 and 0 < N <= self.tp_hidden

    def partial_view(self, M: int, N: int, channel: int = 0) -> torch.Tensor:
        """bf16 [M, N] window of this rank's partial buffer: where a row-parallel GEMM writes when
        the owners pull."""
        return self._bufs[channel, 0][: M * N * 2].view(torch.bfloat16).view(M, N)

    @torch.no_grad()
    def gemm_push(self, qx, sx, qw, sw, bias, dtype, channel: int = 0) -> None:
        """ql_tp_gemm_push: row-parallel W8A8 GEMM whose epilogue delivers every 256-token tile to
        the rank that owns those tokens (GEMM + reduce-scatter in one kernel)."""
        assert qx.dtype == dtype and qw.dtype == dtype and qx.is_contiguous() and qw.is_contiguous()
        M, K = qx.shape
        N = qw.shape[0]
        assert K % 128 == 0 and self.fits_big(M, N)
        _ext._check(self.lib.ql_tp_gemm_push(qx.data_ptr(), sx.data_ptr(), qw.data_ptr(), sw.data_ptr(),
                                             None if bias is None else bias.data_ptr(), M, N, K,
                                             _ext.fmt_of(dtype), channel,
                                             torch.cuda.current_stream().cuda_stream))

    @torch.no_grad()
    def reduce_quant_gather(self, M: int, N: int, dtype, pushed: bool, channel: int = 0):
        """ql_tp_reduce_quant_gather -> (q [M, N], s [M]): views of the channel's symmetric
        buffers, valid until the next prefill-sized collective on that channel."""
        _ext._check(self.lib.ql_tp_reduce_quant_gather(M, N, _ext.fmt_of(dtype), 1 if pushed else 0, channel,
                                                       torch.cuda.current_stream().cuda_stream))
        q = self._bufs[channel, 1][: M * N].view(dtype).view(M, N)
        s = self._bufs[channel, 2][: M * 4].view(torch.float32)
        return q, s

    @torch.no_grad()
    def all_reduce_big(self, M: int, N: int, pushed: bool, device, channel: int = 0,
                       out: torch.Tensor | None = None) -> torch.Tensor:
        """ql_tp_allreduce_bf16 -> bf16 [M, N] (dist.all_reduce semantics), fresh unless ``out``
        (row pitch ``out.stride(0)``) is given."""
        if out is None:
            out = torch.empty((M, N), dtype=torch.bfloat16, device=device)
        assert out.dtype == torch.bfloat16 and out.shape == (M, N) and out.stride(1) == 1
        _ext._check(self.lib.ql_tp_allreduce_bf16(M, N, 1 if pushed else 0, out.data_ptr(), out.stride(0), channel,
                                                  torch.cuda.current_stream().cuda_stream))
        return out


def configure(max_tokens: int | None = None, hidden: int | None = None) -> None:
    """Size the prefill-sized path for the model instead of the generic default (65536 tokens x
    4096 hidden, ~1.4 GB per channel): pass the engine's ``max_num_batched_tokens`` and the model's
    ``hidden_size`` BEFORE the first tensor-parallel forward, identically on every rank.  The memory
    comes out of what the KV cache could otherwise use.  ``QL_TP_TOKENS`` / ``QL_TP_HIDDEN`` in the
    environment still win.  Has no effect once the communicator exists."""
    if max_tokens is not None:
        assert max_tokens > 0
        _state["tokens"] = int(max_tokens)
    if hidden is not None:
        assert hidden > 0 and hidden % 8 == 0
        _state["hidden"] = int(hidden)


def get() -> PeerAllReduce | None:
    """The process-wide communicator, created on first use (never during CUDA-graph capture:
    callers run one eager forward first, as the engine's warm-up does, model_runner.py:91-103).
    Creation is collective: every rank reaches it at the same point of its first forward."""
    if _state["tried"]:
        return _state["comm"]
    if torch.cuda.is_available() and torch.cuda.is_current_stream_capturing():
        return None
    _state["tried"] = True
    if os.environ.get("QL_PEER_AR", "1") == "0":
        return None
    if not (dist.is_available() and dist.is_initialized() and torch.cuda.is_available()):
        return None
    if dist.get_backend() != "nccl" or dist.get_world_size() not in (2, 4, 8):
        return None
    big = os.environ.get("QL_TP_BIG", "1") != "0"
    tokens = int(os.environ.get("QL_TP_TOKENS", str(_state.get("tokens", 65536)))) if big else 0
    hidden = int(os.environ.get("QL_TP_HIDDEN", str(_state.get("hidden", 4096)))) if big else 0
    try:
        _state["comm"] = PeerAllReduce.create(int(os.environ.get("QL_PEER_AR_BYTES", str(1 << 20))),
                                              tokens, hidden)
    except Exception:                        # noqa: BLE001 -- a collective in create() failed: NCCL from here on
        _state["comm"] = None
    return _state["comm"]
