"""Host side of the one-shot NVLink all-reduce (csrc/allreduce.cu, C ABI ``ql_ar_*``).

Used by ``RowParallelLinear.forward`` in place of ``dist.all_reduce(y)`` (reference
nanovllm/layers/linear.py:150-152) when the bf16 output is small (decode-sized
batches, where NCCL's launch + protocol latency exceeds the GEMM).  ``torch.distributed`` is only
the plumbing: it carries the CUDA IPC handles once at start-up.  Larger outputs stay on NCCL.

``QL_PEER_AR=0`` disables it; ``QL_PEER_AR_BYTES`` sets the buffer size (default 1 MiB of payload).
"""
from __future__ import annotations

import ctypes
import os

import torch
import torch.distributed as dist

from nanovllm import _ext

_state = {"tried": False, "comm": None}


class PeerAllReduce:

    def __init__(self, max_bytes: int):
        lib = _ext.load()
        self.world, self.rank = dist.get_world_size(), dist.get_rank()
        hb = lib.ql_ar_handle_bytes()
        mine = ctypes.create_string_buffer(hb)
        _ext._check(lib.ql_ar_init(max_bytes, self.world, self.rank, mine))
        handles = [None] * self.world
        dist.all_gather_object(handles, mine.raw)
        blob = ctypes.create_string_buffer(b"".join(handles), hb * self.world)
        _ext._check(lib.ql_ar_open(blob))
        self.bytes = int(lib.ql_ar_max_bytes())
        torch.cuda.synchronize()
        dist.barrier()                       # every peer has mapped every area before first use

    def fits(self, numel: int) -> bool:
        return numel % 8 == 0 and 0 < numel * 2 <= self.bytes

    @torch.no_grad()
    def all_reduce(self, y: torch.Tensor) -> torch.Tensor:
        """Sum of every rank's y (fp32, rank order, one rounding) -> fresh bf16 tensor."""
        assert y.dtype == torch.bfloat16 and y.is_contiguous() and self.fits(y.numel())
        out = torch.empty_like(y)
        _ext._check(_ext.load().ql_ar_allreduce_bf16(y.data_ptr(), out.data_ptr(), y.numel(),
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
        _ext._check(_ext.load().ql_ar_rmsnorm_quant(
            partial.data_ptr(), residual.data_ptr(), res.data_ptr(), weight.data_ptr(), float(eps), M, K,
            _ext.fmt_of(dtype), q.data_ptr(), s.data_ptr(), torch.cuda.current_stream().cuda_stream))
        return q, s, res


def get() -> PeerAllReduce | None:
    """The process-wide communicator, created on first use (never during CUDA-graph capture:
    callers run one eager forward first, as the engine's warm-up does, model_runner.py:91-103)."""
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
    _state["comm"] = PeerAllReduce(int(os.environ.get("QL_PEER_AR_BYTES", str(1 << 20))))
    return _state["comm"]
