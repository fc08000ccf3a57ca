"""Host logic of the plugin boundary on a CPU box (no GPU): the dispatch of
``ModelRunner.__init__`` (engine/model_runner.py:36-57) and the appliers / holders
(utils/quantization.py:349-401, 463-496) walked over the test-only TinyQwen3.

The product has no CPU arithmetic by design, so -- exactly as tests/test_host_cpu.py does for the
TP layers -- ``nanovllm._ext`` is patched with stand-ins backed by ``oracle/`` (test
infrastructure): what is under test is WHICH module ends up WHERE, with WHICH tensors, and that
the chain of calls is the reference's.  The arithmetic itself is checked on the GPU
(tests/test_model_parity.py).
"""
import numpy as np
import pytest
import torch

from oracle import oracle as orc
from tests import tiny_qwen3 as tq

CFG = dict(hidden=256, inter=512, heads=4, kv=2, hd=64, eps=1e-6, rope_theta=1e6)


@pytest.fixture()
def cpu_ext(monkeypatch):
    from nanovllm import _ext
    calls = {"quant_rows": 0, "quant_groups": 0, "quant_tensor": 0, "linear": 0}

    def quant_rows(x, dtype):
        assert x.dim() == 2
        calls["quant_rows"] += 1
        q, s = orc.quant_rows(orc.t2n(x), orc.fmt_of(dtype))
        return orc.n2t(q, dtype), torch.from_numpy(s)

    def quant_groups(w, g, dtype):
        calls["quant_groups"] += 1
        q, s = orc.quant_groups(orc.t2n(w), g, orc.fmt_of(dtype))
        return orc.n2t(q, dtype), torch.from_numpy(s)

    def quant_tensor(x, dtype):
        calls["quant_tensor"] += 1
        q, sc = orc.quant_tensor(orc.t2n(x), orc.fmt_of(dtype))
        return orc.n2t(q, dtype), torch.tensor(sc, dtype=torch.bfloat16)

    def linear(x, qw, sw, bias, dtype, group_size=0, out=None):
        calls["linear"] += 1
        fmt = orc.fmt_of(dtype)
        x2 = x.reshape(-1, x.shape[-1]).contiguous()
        qx, sx = orc.quant_rows(orc.t2n(x2), fmt)
        b = None if bias is None else orc.t2n(bias)
        if group_size:
            y = orc.gemm_groups(qx, sx, orc.t2n(qw), sw.numpy(), b, group_size, fmt)
        else:
            y = orc.gemm_rows(qx, sx, orc.t2n(qw), sw.numpy(), b, fmt)
        return orc.n2t(y, torch.bfloat16).view(list(x.shape[:-1]) + [qw.shape[0]])

    for name, fn in (("quant_rows", quant_rows), ("quant_groups", quant_groups),
                     ("quant_tensor", quant_tensor), ("linear", linear)):
        monkeypatch.setattr(_ext, name, fn)
    return calls


def _model(factory):
    return tq.TinyQwen3ForCausalLM(2, 128, factory, cfg=CFG, device="cpu", ops="torch")


def _ids():
    return torch.randint(0, 128, (2, 8), generator=torch.Generator().manual_seed(3))


PROJS = (("self_attn", "qkv_proj"), ("self_attn", "o_proj"), ("mlp", "gate_up_proj"), ("mlp", "down_proj"))


@pytest.mark.parametrize("factory", ["tp", "foreign"])
@pytest.mark.parametrize("dtype", [torch.int8, torch.float8_e4m3fn])
@pytest.mark.parametrize("quant_type,group", [("per_row", 128), (None, 128), ("per_group", 64),
                                              ("per_group", 128), ("per_tensor", 128)])
def test_dispatch_selects_the_reference_branch(cpu_ext, factory, dtype, quant_type, group):
    from nanovllm.layers.linear import LinearBase
    from nanovllm.utils import quantization as Q
    fac = tq.tp_linears if factory == "tp" else tq.foreign_linears
    m = _model(fac)
    Q.apply_quantization(m, dtype, quant_type, group)
    eff = quant_type or "per_row"                      # model_runner.py:55-57: dtype alone => per-row
    holder = {"per_row": Q.QLinear, "per_group": Q.QGroupLinear, "per_tensor": Q.QTensorLinear}[eff]
    n_lin = 0
    for layer in m.model.layers:
        for parent, name in PROJS:
            lin = getattr(getattr(layer, parent), name)
            n_lin += 1
            if factory == "tp":
                assert isinstance(lin, LinearBase) and lin.weight is None and lin.quant_type == eff
            else:
                assert type(lin) is holder
            assert lin.qw.dtype == dtype
            N, K = lin.qw.shape
            if eff == "per_row":
                assert lin.sw.shape == (N,) and lin.sw.dtype == torch.float32
            elif eff == "per_group":
                assert lin.sw.shape == (N, K // group) and lin.group_size == group
            else:
                assert lin.sw.dim() == 0 and lin.sw.dtype == torch.bfloat16   # reference :166-191
    key = {"per_row": "quant_rows", "per_group": "quant_groups", "per_tensor": "quant_tensor"}[eff]
    assert cpu_ext[key] == n_lin == 8
    # the untouched neighbours (quantization.py:463-496 never visits them)
    assert m.lm_head.weight.dtype == torch.bfloat16 and m.model.embed_tokens.weight.dtype == torch.bfloat16
    # forward goes through the quantized path once per projection and equals the oracle chain
    logits = m(_ids())
    assert cpu_ext["linear"] == n_lin
    ref = _model(tq.foreign_linears)
    ref.swap_linears(lambda w: tq.OracleLinear(w, dtype, eff, group, "w8a8"))
    assert torch.equal(logits, ref(_ids()))


def test_dispatch_corners(cpu_ext):
    from nanovllm.utils import quantization as Q
    m = _model(tq.tp_linears)
    Q.apply_quantization(m, torch.bfloat16, None, 128)         # nothing selected -> bf16 model
    assert all(not layer.mlp.down_proj.is_quantized for layer in m.model.layers)
    for bad in ("per_banana", "smoothquant"):
        with pytest.raises(ValueError):
            Q.apply_quantization(m, torch.int8, bad, 128)
    with pytest.raises(AssertionError):                        # reference :353, :371, :389
        Q.apply_quantization(m, torch.float16, "per_row", 128)
    assert sum(cpu_ext.values()) == 0
    lin = tq.ForeignLinear(256, 128).to(torch.bfloat16)
    lin.tp_size = 2
    with pytest.raises(AssertionError):                        # reference :360 "Not support TP."
        Q.QLinear.from_linear(lin, torch.int8)


def test_holder_bias_and_3d_input(cpu_ext):
    """QLinear keeps the original bias Parameter and accepts [1, L, K] (test_ppl.py's shape)."""
    from nanovllm.utils import quantization as Q
    g = torch.Generator().manual_seed(5)
    lin = tq.ForeignLinear(256, 128).to(torch.bfloat16)
    lin.weight.data.copy_((torch.randn(128, 256, generator=g) * 0.05).to(torch.bfloat16))
    lin.bias = torch.nn.Parameter((torch.randn(128, generator=g)).to(torch.bfloat16))
    x = torch.randn(1, 6, 256, generator=g).to(torch.bfloat16)
    for holder in (Q.QLinear.from_linear(lin, torch.int8), Q.QTensorLinear.from_linear(lin, torch.int8),
                   Q.QGroupLinear.from_linear(lin, torch.int8, 128)):
        assert holder.bias is lin.bias
        y = holder(x)
        assert y.shape == (1, 6, 128) and y.dtype == torch.bfloat16
        ref = torch.nn.functional.linear(x.float(), lin.weight.float(), lin.bias.float())
        assert float((y.float() - ref).abs().max() / ref.abs().max()) < 0.08
