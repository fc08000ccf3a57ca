"""RMSNorm with the per-token quantiser fused in.

Mirrors the reference's nanovllm/layers/layernorm.py (class name, constructor, ``forward(x,
residual=None)`` returning ``x`` or ``(x, residual)``).  The reference compiles two torch
functions (rms_forward :16-25, add_rms_forward :27-40); here both run in ONE sm_100a kernel
(``ql_rmsnorm_quant``), and ``forward_quant`` additionally hands the following quantized linear
its input already quantised -- ``QuantizedActivation(q, s)`` -- so the bf16 normalised tensor
never goes to HBM and the linear skips its own ``per_row_quant`` launch.

No torch / CPU fallback: CUDA tensors on a B200 only.
"""
from __future__ import annotations

import torch
from torch import nn

from nanovllm import _ext


class QuantizedActivation:
    """Per-token quantised activation: ``q`` [M, K] int8 / float8_e4m3fn, ``s`` [M] fp32 --
    exactly what ``per_row_quant(x, dtype)`` returns (utils/quantization.py:147-163).  The
    quantized linear layers accept it in place of the bf16 tensor."""

    __slots__ = ("q", "s")

    def __init__(self, q: torch.Tensor, s: torch.Tensor):
        self.q, self.s = q, s

    @property
    def dtype(self) -> torch.dtype:
        return self.q.dtype

    @property
    def shape(self):
        return self.q.shape


class RMSNorm(nn.Module):

    def __init__(self, hidden_size: int, eps: float = 1e-6) -> None:
        super().__init__()
        self.eps = eps
        self.weight = nn.Parameter(torch.ones(hidden_size))

    def _run(self, x, residual, dtype, want_bf16):
        K = x.shape[-1]
        x2 = x.reshape(-1, K)
        if x2.stride(-1) != 1 or (x2.shape[0] > 1 and x2.stride(0) % 8 != 0) or x2.data_ptr() % 16:
            x2 = x2.contiguous()                  # the kernel needs 16-byte aligned, K-contiguous rows
        res_out = None
        if residual is not None:
            # add_rms_forward returns a NEW residual tensor (:34): the kernel writes x + residual
            # into a fresh buffer, the caller's tensor is not modified
            residual = residual.reshape(-1, K).contiguous()
            res_out = torch.empty_like(residual)
        q, s, h = _ext.rmsnorm_quant(x2, self.weight.data, self.eps, residual, dtype, want_bf16,
                                     residual_out=res_out)
        return q, s, h, res_out

    def forward(self, x: torch.Tensor, residual: torch.Tensor | None = None):
        """bf16 in, bf16 out -- the reference's contract (:42-50)."""
        _, _, h, res = self._run(x, residual, None, True)
        h = h.view(x.shape)
        return h if residual is None else (h, res.view(x.shape))

    def forward_quant(self, x: torch.Tensor, residual: torch.Tensor | None, dtype: torch.dtype,
                      partial: bool = False):
        """As ``forward`` but returns the normalised activation quantised per token for the next
        quantized linear: ``QuantizedActivation`` or ``(QuantizedActivation, residual)``.

        ``partial=True`` (tensor parallel): ``x`` is this rank's UN-REDUCED row-parallel output
        (``RowParallelLinear.forward_partial``); the all-reduce runs inside the same kernel
        (``ql_ar_rmsnorm_quant``) when the batch is decode-sized, else NCCL + the plain kernel."""
        if partial:
            import torch.distributed as dist
            from nanovllm import _comm
            assert residual is not None
            comm = _comm.get()
            x2 = x.reshape(-1, x.shape[-1])
            if comm is not None and x2.is_contiguous() and comm.fits_fused(*x2.shape):
                q, s, res = comm.all_reduce_rmsnorm_quant(x2, residual.reshape(x2.shape).contiguous(),
                                                          self.weight.data, self.eps, dtype)
                return QuantizedActivation(q, s), res.view(x.shape)
            x = x.clone()
            dist.all_reduce(x)
        q, s, _, res = self._run(x, residual, dtype, False)
        qa = QuantizedActivation(q, s)
        return qa if residual is None else (qa, res.view(x.shape))
