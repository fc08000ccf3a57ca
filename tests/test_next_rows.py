"""quantized LM head, serialised quantized checkpoints."""
import os

import numpy as np
import pytest
import torch

from oracle import oracle as orc


def _make_layers(tp_rank, tp_size, monkeypatch):
    """The four TP layer classes of a (small) Qwen3 layer as rank `tp_rank` of `tp_size`, on CPU."""
    import nanovllm.layers.linear as L
    monkeypatch.setattr(L, "_tp_rank_size", lambda: (tp_rank, tp_size))
    return {"qkv": L.QKVParallelLinear(256, 64, 4, 2, bias=True),          # N = (4 + 2*2) * 64 = 512
            "o": L.RowParallelLinear(256, 256),
            "gate_up": L.MergedColumnParallelLinear(256, [512, 512]),
            "down": L.RowParallelLinear(512, 256, bias=True)}


def _fake_quantize(layer, seed, quant_type, g):
    """Quantized state from the CPU oracle (the product has no CPU quantiser)."""
    n, k = layer.weight.shape
    w = (torch.randn(n, k, generator=torch.Generator().manual_seed(seed)) * 0.02).to(torch.bfloat16)
    if quant_type == "per_row":
        q, s = orc.quant_rows(orc.t2n(w), orc.INT8)
        layer.group_size = 0
    else:
        q, s = orc.quant_groups(orc.t2n(w), g, orc.INT8)
        layer.group_size = g
    layer.qw, layer.sw = orc.n2t(q, torch.int8), torch.from_numpy(np.ascontiguousarray(s))
    layer._sw_kernel, layer.qdtype, layer.quant_type = layer.sw, torch.int8, quant_type
    if layer.bias is not None:
        layer.bias.data.normal_()
    layer.weight = None


@pytest.mark.parametrize("quant_type", ["per_row", "per_group"])
def test_checkpoint_sharding_matches_weight_loaders(tmp_path, monkeypatch, quant_type):
    """TP=1 checkpoint -> TP=2 ranks: the code / scale rows each rank receives are exactly the rows
    the reference's weight_loaders (linear.py:54-153) would hand it."""
    from nanovllm.utils.qcheckpoint import load_quantized, save_quantized
    full = _make_layers(0, 1, monkeypatch)
    for i, (name, layer) in enumerate(full.items()):
        _fake_quantize(layer, i, quant_type, 128)
    path = os.path.join(tmp_path, "q.safetensors")
    save_quantized(full, path)
    # round trip at TP=1
    again = _make_layers(0, 1, monkeypatch)
    load_quantized(again, path, device="cpu")
    for name in full:
        assert torch.equal(again[name].qw, full[name].qw) and torch.equal(again[name].sw, full[name].sw)
        assert again[name].weight is None and again[name].quant_type == quant_type
    for rank in range(2):
        shard = _make_layers(rank, 2, monkeypatch)
        # which rows / columns does this rank own?  ask the bf16 weight_loaders with index tensors
        owners = {}
        for name, layer in shard.items():
            n, k = full[name].qw.shape
            row_id = torch.arange(n, dtype=torch.float32)[:, None].expand(n, k).contiguous()
            col_id = torch.arange(k, dtype=torch.float32)[None, :].expand(n, k).contiguous()
            got = {}
            for tag, idx in (("row", row_id), ("col", col_id)):
                layer.weight.data = torch.zeros_like(layer.weight.data, dtype=torch.float32)
                if name == "qkv":
                    q_rows, kv_rows = 4 * 64, 2 * 64
                    for sid, a, b in (("q", 0, q_rows), ("k", q_rows, q_rows + kv_rows),
                                      ("v", q_rows + kv_rows, q_rows + 2 * kv_rows)):
                        layer.weight_loader(layer.weight, idx[a:b], sid)
                elif name == "gate_up":
                    for sid in (0, 1):
                        layer.weight_loader(layer.weight, idx[sid * 512:(sid + 1) * 512], sid)
                else:
                    layer.weight_loader(layer.weight, idx)
                got[tag] = layer.weight.data.clone()
            owners[name] = (got["row"][:, 0].long(), got["col"][0, :].long())
        load_quantized(shard, path, device="cpu")
        for name, layer in shard.items():
            rows, cols = owners[name]
            assert torch.equal(layer.qw, full[name].qw[rows][:, cols]), name
            if quant_type == "per_row":
                assert torch.equal(layer.sw, full[name].sw[rows]), name
            else:
                gcols = torch.unique(cols // 128)
                assert torch.equal(layer.sw, full[name].sw[rows][:, gcols]), name
            if layer.bias is not None:
                want = full[name].bias.data[rows] if layer.tp_dim == 0 else full[name].bias.data
                assert torch.equal(layer.bias.data.float(), want.float())


# ------------------------------------------------------------------ GPU

@pytest.mark.gpu
@pytest.mark.parametrize("dn", ["i8", "f8"])
@pytest.mark.parametrize("M", [1, 16, 200])
def test_quantized_lm_head_vocab_shape(dn, M):
    """Qwen3 vocabulary (151 936 x 1024): logits of the quantized head vs the oracle on the first,
    middle and last vocabulary rows (per-row outputs are independent per row)."""
    from nanovllm.layers.embed_head import ParallelLMHead
    dt = {"i8": torch.int8, "f8": torch.float8_e4m3fn}[dn]
    fmt = {"i8": orc.INT8, "f8": orc.FP8}[dn]
    V, H = 151936, 1024
    prev = torch.get_default_dtype()
    torch.set_default_dtype(torch.bfloat16)
    with torch.device("cuda"):
        head = ParallelLMHead(V, H)
    torch.set_default_dtype(prev)
    head.weight.data.normal_(0, 0.02, generator=torch.Generator(device="cuda").manual_seed(5))
    w_cpu = head.weight.data.cpu()
    head.quantize_(dt, keep_weight=True)
    assert head.weight is not None and head.qw.shape == (V, H)
    from nanovllm.utils.context import reset_context, set_context
    x = torch.randn(M + 3, H, generator=torch.Generator().manual_seed(M)).to(torch.bfloat16)
    # the reference's calling convention (embed_head.py:56-66): forward(x), last-token selection
    # from the prefill context.  M sequences; the first is 4 tokens long, the others 1 token each,
    # so cu_seqlens_q[1:] - 1 = 3, 4, ..., M + 2.
    cu = torch.cat([torch.tensor([0]), torch.arange(4, M + 4)]).to(torch.int32).cuda()
    set_context(True, cu_seqlens_q=cu)
    try:
        logits = head(x.cuda())
    finally:
        reset_context()
    assert logits.shape == (M, V) and logits.dtype == torch.bfloat16
    n_dec = min(M, 2)
    dec = head(x[3:3 + n_dec].cuda())                         # decode: no selection (context.is_prefill False)
    assert dec.shape == (n_dec, V)
    assert float((dec.float() - logits[:n_dec].float()).abs().max()) <= 0.02 * float(logits.float().abs().max())
    idx = np.r_[0:300, V // 2:V // 2 + 300, V - 300:V]
    qwo, swo = orc.quant_rows(orc.t2n(w_cpu[idx]), fmt)
    assert np.array_equal(orc.t2n(head.qw[idx]), qwo) and np.array_equal(orc.t2n(head.sw[idx]), swo)
    qxo, sxo = orc.quant_rows(orc.t2n(x[3:]), fmt)
    yo = orc.gemm_rows(qxo, sxo, qwo, swo, None, fmt)
    got = orc.t2n(logits[:, idx])
    if dn == "i8":
        assert np.array_equal(got, yo)                        # bit-exact
    else:
        err = np.abs(orc.bf16_decode(got) - orc.bf16_decode(yo)).max()
        assert err <= 2.0 ** (np.floor(np.log2(np.abs(orc.bf16_decode(yo)).max())) - 7)   # 1 bf16 ulp of max|y|


@pytest.mark.gpu
@pytest.mark.parametrize("quant_type", ["per_row", "per_group"])
def test_checkpoint_roundtrip_forward_identical(tmp_path, quant_type):
    from nanovllm.layers.linear import MergedColumnParallelLinear, RowParallelLinear
    from nanovllm.utils.qcheckpoint import load_quantized, save_quantized

    def build():
        prev = torch.get_default_dtype()
        torch.set_default_dtype(torch.bfloat16)
        with torch.device("cuda"):
            layers = {"mlp.gate_up_proj": MergedColumnParallelLinear(1024, [3072, 3072]),
                      "mlp.down_proj": RowParallelLinear(3072, 1024, bias=True)}
        torch.set_default_dtype(prev)
        return layers
    src = build()
    for i, layer in enumerate(src.values()):
        layer.weight.data.normal_(0, 0.02, generator=torch.Generator(device="cuda").manual_seed(i))
        if layer.bias is not None:
            layer.bias.data.normal_()
        layer.quantize_(torch.float8_e4m3fn, quant_type, 128)
    path = os.path.join(tmp_path, "q.safetensors")
    save_quantized(src, path)
    dst = build()
    load_quantized(dst, path)
    x = torch.randn(40, 1024, device="cuda").to(torch.bfloat16)
    for name in src:
        xin = x if "gate_up" in name else torch.randn(40, 3072, device="cuda").to(torch.bfloat16)
        assert dst[name].weight is None
        assert torch.equal(src[name](xin), dst[name](xin)), name


def test_checkpoint_rejects_mismatched_file(tmp_path, monkeypatch):
    """A checkpoint written for other shapes, or one without the bias a layer has, must not load."""
    import pytest
    from nanovllm.layers.linear import ReplicatedLinear
    from nanovllm.utils import qcheckpoint
    from safetensors.torch import save_file
    import json
    N, K = 16, 128
    good = {"l.qweight": torch.zeros(N, K, dtype=torch.int8).view(torch.uint8), "l.scale": torch.ones(N)}
    meta = {"qlinear_b200": json.dumps({"l": "int8,per_row,0"})}
    prev = torch.get_default_dtype()
    torch.set_default_dtype(torch.bfloat16)
    try:
        layer = ReplicatedLinear(K, N, bias=True)
        other = ReplicatedLinear(K, 2 * N)
    finally:
        torch.set_default_dtype(prev)
    p = str(tmp_path / "q.safetensors")
    save_file(good, p, metadata=meta)
    with pytest.raises(ValueError, match="bias"):
        qcheckpoint.load_quantized({"l": layer}, p, device="cpu")
    with pytest.raises(ValueError, match="qweight"):
        qcheckpoint.load_quantized({"l": other}, p, device="cpu")
    bad = dict(good)
    bad["l.scale"] = torch.ones(N + 1)
    save_file(bad, p, metadata=meta)
    layer2 = ReplicatedLinear(K, N)
    with pytest.raises(ValueError, match="scale"):
        qcheckpoint.load_quantized({"l": layer2}, p, device="cpu")
