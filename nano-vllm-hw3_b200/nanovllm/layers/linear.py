"""Tensor-parallel linear layers with the quantized path INSIDE them.

Mirrors the reference's nanovllm/layers/linear.py (class names, constructor signatures,
``weight_loader(param, tensor[, shard_id])`` contracts, ``tp_dim/tp_rank/tp_size``,
``weight.weight_loader`` attribute used by utils/loader.py:21-23):

    LinearBase :12-34, ReplicatedLinear :37-51, ColumnParallelLinear :54-73,
    MergedColumnParallelLinear :76-93, QKVParallelLinear :96-128, RowParallelLinear :131-153

What is new: ``quantize_(dtype, quant_type, group_size)`` converts the already-sharded bf16
``weight`` into (``qw``, ``sw``) per TP shard, and ``forward`` then runs the sm_100a W8A8
kernels (``nanovllm._ext.linear``).  The reference instead swaps the module for a
``QLinear`` that asserts ``tp_size == 1`` (utils/quantization.py:358-361); here
quantization composes with tensor parallelism: column-parallel shards N (row-local scales
commute with the split), row-parallel shards K (each rank quantises its own K-slice, adds
bias on rank 0 only and all-reduces the bf16 partial sums, as linear.py:150-152).
"""
from __future__ import annotations

import contextlib
import os

import torch
import torch.distributed as dist
import torch.nn.functional as F
from torch import nn

from nanovllm import _ext


_tp_channel = [0]


@contextlib.contextmanager
def tp_channel(channel: int):
    """Route the prefill-sized NVLink collectives of the row-parallel linears called inside the
    block through ``channel`` (0 .. ql_tp_channels()-1; independent buffers, flags and call counters,
    include/qlinear_b200.h).  A caller that runs two token micro-batches on two CUDA streams gives
    each its own channel; every rank must use the same channel for the same micro-batch.

    KNOWN HAZARD: two channels in flight at once can hang at TP=8 (a cross-GPU placement cycle
    between one channel's spinning reduce CTAs and the other's persistent GEMM; see
    include/qlinear_b200.h, ql_tp_channels).  The mode is therefore refused -- RuntimeError --
    unless ``QL_TP_ALLOW_CONCURRENT=1`` is set in the environment."""
    if os.environ.get("QL_TP_ALLOW_CONCURRENT", "")[:1] != "1":
        raise RuntimeError(
            "tp_channel: the two-channel tensor-parallel pipeline has a known cross-GPU ordering hazard "
            "(can hang at TP=8); set QL_TP_ALLOW_CONCURRENT=1 to opt in")
    prev = _tp_channel[0]
    _ext._check(_ext.load().ql_tp_set_concurrent(1))   # reduce kernels release their stream late from here on
    _tp_channel[0] = int(channel)
    try:
        yield
    finally:
        _tp_channel[0] = prev


def divide(numerator, denominator):
    assert numerator % denominator == 0
    return numerator // denominator


def _tp_rank_size():
    if dist.is_available() and dist.is_initialized():
        return dist.get_rank(), dist.get_world_size()
    return 0, 1


class LinearBase(nn.Module):

    def __init__(self, input_size: int, output_size: int, bias: bool = False,
                 tp_dim: int | None = None):
        super().__init__()
        self.tp_dim = tp_dim
        self.tp_rank, self.tp_size = _tp_rank_size()
        self.weight = nn.Parameter(torch.empty(output_size, input_size))
        self.weight.weight_loader = self.weight_loader
        if bias:
            self.bias = nn.Parameter(torch.empty(output_size))
            self.bias.weight_loader = self.weight_loader
        else:
            self.register_parameter("bias", None)
        # quantized state (plain attributes, like the reference's QLinear.qw/.sw: never
        # Parameters or buffers, never serialised)
        self.qdtype: torch.dtype | None = None
        self.quant_type: str | None = None
        self.group_size: int = 0
        self.qw: torch.Tensor | None = None
        self.sw: torch.Tensor | None = None
        self._sw_kernel: torch.Tensor | None = None

    # -------------------------------------------------------------- quantization
    @torch.no_grad()
    def quantize_(self, dtype: torch.dtype, quant_type: str | None = "per_row",
                  group_size: int = 128) -> "LinearBase":
        """Quantise this rank's weight shard in place and drop the bf16 copy.

        ``quant_type``: "per_row" (default rule of model_runner.py:55-57), "per_group",
        "per_tensor".  Layout afterwards is the reference's: ``qw [N,K]`` in ``dtype``,
        ``sw`` fp32 ``[N]`` / ``[N, K/group_size]`` / 0-dim bf16 (per-tensor).
        """
        from nanovllm.utils import quantization as Q
        assert dtype in [torch.float8_e4m3fn, torch.int8]
        assert self.weight is not None, "already quantized"
        quant_type = quant_type or "per_row"
        w = self.weight.data
        if quant_type == "per_row":
            self.qw, self.sw = Q.per_row_quant(w, dtype)
            self._sw_kernel, self.group_size = self.sw, 0
        elif quant_type == "per_group":
            self.qw, self.sw = Q.per_group_quant(w, group_size, dtype)
            self._sw_kernel, self.group_size = self.sw, group_size
        elif quant_type == "per_tensor":
            self.qw, self.sw = Q.per_tensor_quant(w, dtype)
            self._sw_kernel = self.sw.float().expand(w.size(0)).contiguous()
            self.group_size = 0
        else:
            raise ValueError(f"Unknown quant_type: {quant_type}")
        self.qdtype, self.quant_type = dtype, quant_type
        self.weight = None
        return self

    @property
    def is_quantized(self) -> bool:
        return self.qw is not None

    def _linear(self, x, bias: torch.Tensor | None, out: torch.Tensor | None = None) -> torch.Tensor:
        if self.qw is None:
            return F.linear(x, self.weight, bias)
        b = None if bias is None else bias.data
        if hasattr(x, "q") and hasattr(x, "s"):
            # QuantizedActivation from a fused producer (layers/layernorm.py, layers/activation.py):
            # this linear's per_row_quant launch is skipped
            assert x.q.dtype == self.qdtype, "producer and linear must use the same 8-bit format"
            return _ext.gemm_w8a8(x.q, x.s, self.qw, self._sw_kernel, b, self.qdtype, self.group_size,
                                  out=out)
        return _ext.linear(x, self.qw, self._sw_kernel, b, self.qdtype, self.group_size, out=out)

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        raise NotImplementedError


class ReplicatedLinear(LinearBase):

    def __init__(self, input_size: int, output_size: int, bias: bool = False):
        super().__init__(input_size, output_size, bias)

    def weight_loader(self, param: nn.Parameter, loaded_weight: torch.Tensor):
        param.data.copy_(loaded_weight)

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        return self._linear(x, self.bias)


class ColumnParallelLinear(LinearBase):

    def __init__(self, input_size: int, output_size: int, bias: bool = False):
        _, tp_size = _tp_rank_size()
        super().__init__(input_size, divide(output_size, tp_size), bias, 0)

    def weight_loader(self, param: nn.Parameter, loaded_weight: torch.Tensor):
        shard = param.data.size(self.tp_dim)
        param.data.copy_(loaded_weight.narrow(self.tp_dim, self.tp_rank * shard, shard))

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        return self._linear(x, self.bias)


class MergedColumnParallelLinear(ColumnParallelLinear):

    def __init__(self, input_size: int, output_sizes: list[int], bias: bool = False):
        self.output_sizes = output_sizes
        super().__init__(input_size, sum(output_sizes), bias)

    def weight_loader(self, param: nn.Parameter, loaded_weight: torch.Tensor, loaded_shard_id: int):
        offset = sum(self.output_sizes[:loaded_shard_id]) // self.tp_size
        size = self.output_sizes[loaded_shard_id] // self.tp_size
        dst = param.data.narrow(self.tp_dim, offset, size)
        dst.copy_(loaded_weight.chunk(self.tp_size, self.tp_dim)[self.tp_rank])


class QKVParallelLinear(ColumnParallelLinear):

    def __init__(self, hidden_size: int, head_size: int, total_num_heads: int,
                 total_num_kv_heads: int | None = None, bias: bool = False):
        _, tp_size = _tp_rank_size()
        total_num_kv_heads = total_num_kv_heads or total_num_heads
        self.head_size = head_size
        self.num_heads = divide(total_num_heads, tp_size)
        self.num_kv_heads = divide(total_num_kv_heads, tp_size)
        super().__init__(hidden_size, (total_num_heads + 2 * total_num_kv_heads) * head_size, bias)

    def weight_loader(self, param: nn.Parameter, loaded_weight: torch.Tensor, loaded_shard_id: str):
        assert loaded_shard_id in ["q", "k", "v"]
        q_rows = self.num_heads * self.head_size
        kv_rows = self.num_kv_heads * self.head_size
        offset, size = {"q": (0, q_rows), "k": (q_rows, kv_rows),
                        "v": (q_rows + kv_rows, kv_rows)}[loaded_shard_id]
        dst = param.data.narrow(self.tp_dim, offset, size)
        dst.copy_(loaded_weight.chunk(self.tp_size, self.tp_dim)[self.tp_rank])


class RowParallelLinear(LinearBase):

    def __init__(self, input_size: int, output_size: int, bias: bool = False):
        _, tp_size = _tp_rank_size()
        super().__init__(divide(input_size, tp_size), output_size, bias, 1)

    def weight_loader(self, param: nn.Parameter, loaded_weight: torch.Tensor):
        if param.data.dim() == 1:          # bias [N] is replicated, not sharded along K
            param.data.copy_(loaded_weight)
            return
        shard = param.data.size(self.tp_dim)
        param.data.copy_(loaded_weight.narrow(self.tp_dim, self.tp_rank * shard, shard))

    def forward_partial(self, x) -> torch.Tensor:
        """This rank's un-reduced output (bias on rank 0 only): the input of the fused
        all-reduce + RMSNorm + quant (`RMSNorm.forward_quant(..., partial=True)`)."""
        return self._linear(x, self.bias if self.tp_rank == 0 else None)

    def _comm(self):
        if self.tp_size == 1 or self.qw is None:
            return None
        from nanovllm import _comm
        return _comm.get()

    def _big_partial(self, comm, x, bias):
        """Prefill-sized output: leave this rank's partial sums where the token owners can reduce
        them.  Returns (M, N, pushed).  Per-row scales: the CTA-pair GEMM pushes every 256-token
        tile into its owner's receive region from its epilogue (GEMM + reduce-scatter in one
        kernel); otherwise (per-group scales, QL_TP_PUSH=0) the GEMM writes into this rank's
        window of the symmetric area and the owners pull."""
        b = None if bias is None else bias.data
        N = self.qw.shape[0]
        if comm.push and self.group_size == 0:
            if hasattr(x, "q") and hasattr(x, "s"):
                qx, sx = x.q, x.s
            else:
                qx, sx = _ext.quant_rows(x.reshape(-1, x.shape[-1]), self.qdtype)
            comm.gemm_push(qx, sx, self.qw, self._sw_kernel, b, self.qdtype, _tp_channel[0])
            return qx.shape[0], N, True
        M = x.q.shape[0] if hasattr(x, "q") else x.reshape(-1, x.shape[-1]).shape[0]
        self._linear(x, bias, out=comm.partial_view(M, N, _tp_channel[0]))
        return M, N, False

    def forward_quant(self, x):
        """Row-parallel linear whose all-reduced output is handed to the NEXT quantized linear
        already quantised per token (``QuantizedActivation``): under tensor parallelism the
        reduction is a reduce-scatter to the token owner, the owner quantises its tokens only, and
        codes + scales are all-gathered -- instead of `dist.all_reduce` (linear.py:150-152)
        followed by `per_row_quant` over all tokens on every rank (utils/quantization.py:323).
        Bit-identical to all-reduce (fp32, rank order) + per_row_quant."""
        from nanovllm.layers.layernorm import QuantizedActivation
        assert self.qw is not None, "forward_quant needs a quantized layer"
        bias = self.bias if self.tp_rank == 0 else None
        comm = self._comm()
        M = x.q.shape[0] if hasattr(x, "q") else x.reshape(-1, x.shape[-1]).shape[0]
        N = self.qw.shape[0]
        if comm is not None and comm.fits_big(M, N) and not comm.fits(M * N):
            M, N, pushed = self._big_partial(comm, x, bias)
            return QuantizedActivation(*comm.reduce_quant_gather(M, N, self.qdtype, pushed, _tp_channel[0]))
        y = self.forward(x)
        return QuantizedActivation(*_ext.quant_rows(y.reshape(-1, N), self.qdtype))

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        bias = self.bias if self.tp_rank == 0 else None
        if self.tp_size == 1:
            return self._linear(x, bias)
        comm = self._comm()
        if comm is not None:
            M = x.q.shape[0] if hasattr(x, "q") else x.reshape(-1, x.shape[-1]).shape[0]
            N = self.qw.shape[0]
            lead = list(x.q.shape[:-1] if hasattr(x, "q") else x.shape[:-1])
            if comm.fits(M * N):
                # decode-sized output: one push-protocol kernel over NVLink peer memory
                # (csrc/allreduce.cu) instead of the NCCL call
                y = self._linear(x, bias)
                if y.is_contiguous():
                    return comm.all_reduce(y)
                dist.all_reduce(y)
                return y
            if comm.fits_big(M, N):
                # prefill-sized output: reduce-scatter (from the GEMM epilogue) + all-gather
                M, N, pushed = self._big_partial(comm, x, bias)
                dev = x.q.device if hasattr(x, "q") else x.device
                return comm.all_reduce_big(M, N, pushed, dev, _tp_channel[0]).view(lead + [N])
        y = self._linear(x, bias)
        dist.all_reduce(y)                               # reference linear.py:150-152
        return y
