"""CPU-side tests (no GPU): the C-ABI library loads and exports every symbol the header
declares, argument validation happens before any device work, the product path refuses to
run without a B200, and the tensor-parallel host logic (weight_loader sharding, rank-0 bias,
all-reduce after the row-parallel linear; reference nanovllm/layers/linear.py:54-153) is
right at world_size 2 over gloo.

The TP test needs arithmetic on a CPU box, where the product has none by design: it patches
``nanovllm._ext`` with stand-ins backed by ``oracle/`` (test infrastructure) so that only
the HOST logic of the layers is under test.
"""
import ctypes
import os
import re
import socket
import sys

import numpy as np
import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "qlinear_b200.h")


@pytest.fixture(scope="module")
def lib():
    sys.path.insert(0, ROOT)
    import __graft_entry__ as ge
    ge.build()
    from nanovllm import _ext
    return _ext.load()


def declared_symbols():
    text = open(HEADER).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"QL_API\s+[\w\s\*]+?\b(ql_\w+)\s*\(", text)))


def test_header_declares_the_documented_surface():
    names = declared_symbols()
    for must in ["ql_version", "ql_status_string", "ql_quant_rows", "ql_quant_groups", "ql_gemm_w8a8",
                 "ql_gemv_w8a8", "ql_linear", "ql_workspace_bytes"]:
        assert must in names


def test_library_exports_every_declared_symbol(lib):
    from nanovllm import _ext
    raw = ctypes.CDLL(_ext.LIB_PATH)
    for name in declared_symbols():
        assert hasattr(raw, name), f"{name} declared in include/qlinear_b200.h but not exported"
        assert name in _ext.SIGNATURES, f"{name} has no ctypes signature in nanovllm/_ext.py"
    assert set(_ext.SIGNATURES) == set(declared_symbols())


def test_version_and_status_strings(lib):
    from nanovllm import _ext
    m = re.search(r"#define QL_ABI_VERSION (\d+)", open(HEADER).read())
    assert lib.ql_version() == int(m.group(1)) == _ext.QL_ABI_VERSION
    seen = set()
    for code in range(0, 8):
        s = lib.ql_status_string(code).decode()
        assert s and s != "unknown status"
        seen.add(s)
    assert len(seen) == 8
    assert lib.ql_status_string(99).decode() == "unknown status"


def test_workspace_bytes(lib):
    assert lib.ql_workspace_bytes(0, 1024) == 0
    # decode regime: optional (the fused kernels run without it) but sized for the two-launch path
    assert lib.ql_workspace_bytes(16, 4096) >= 16 * 4096 + 16 * 4
    n = lib.ql_workspace_bytes(17, 1024)
    assert n >= 17 * 1024 + 17 * 4 and n % 256 == 0
    assert lib.ql_workspace_bytes(4096, 3072) >= 4096 * 3072 + 4096 * 4


def test_python_workspace_size_mirrors_the_library(lib):
    from nanovllm import _ext
    for M, K in [(0, 1024), (1, 128), (16, 4096), (17, 1024), (4096, 3072), (33, 12288), (8192, 1024)]:
        assert _ext.workspace_bytes(M, K) == lib.ql_workspace_bytes(M, K), (M, K)


def test_validation_precedes_device_work(lib):
    """Bad arguments are rejected with their own code on a box without any GPU."""
    fake = ctypes.c_void_p(0x1000)
    # K % 128 != 0 (reference quantization.py:328)
    assert lib.ql_linear(fake, 100, fake, fake, None, fake, 8, 4, 8, 100, 0, 0, None, 0, None) == 2
    # bad format (reference :151)
    assert lib.ql_quant_rows(fake, 4, 128, 128, 7, fake, fake, None) == 1
    # K % group != 0 (reference :205)
    assert lib.ql_quant_groups(fake, 4, 256, 96, 0, fake, fake, None) == 2
    # GEMV is M <= 16 only
    assert lib.ql_gemv_w8a8(fake, 128, fake, fake, None, fake, 8, 17, 8, 128, 0, 0, None) == 2
    # NULL pointers
    assert lib.ql_gemm_w8a8(None, fake, fake, fake, None, fake, 8, 4, 8, 128, 0, 0, None) == 7
    # misaligned weight pointer
    assert lib.ql_gemm_w8a8(fake, fake, ctypes.c_void_p(0x1008), fake, None, fake, 8, 4, 8, 128, 0, 0, None) == 3
    # empty problems are OK without touching the device
    assert lib.ql_linear(fake, 128, fake, fake, None, fake, 8, 0, 8, 128, 0, 0, None, 0, None) == 0


def test_two_channel_tp_mode_needs_an_explicit_opt_in(lib, monkeypatch):
    """the two-channel pipeline can hang at TP=8; it is refused unless opted into."""
    from nanovllm.layers import linear as L
    monkeypatch.delenv("QL_TP_ALLOW_CONCURRENT", raising=False)
    assert lib.ql_tp_set_concurrent(1) == 9
    assert lib.ql_tp_reduce_quant_gather(512, 1024, 0, 0, 1, None) == 9
    assert lib.ql_tp_gemm_push(None, None, None, None, None, 512, 1024, 1024, 0, 1, None) == 9
    assert "QL_TP_ALLOW_CONCURRENT" in lib.ql_status_string(9).decode()
    with pytest.raises(RuntimeError, match="QL_TP_ALLOW_CONCURRENT"):
        with L.tp_channel(1):
            pass
    assert lib.ql_tp_set_concurrent(0) == 0
    monkeypatch.setenv("QL_TP_ALLOW_CONCURRENT", "1")
    assert lib.ql_tp_set_concurrent(1) == 0
    with L.tp_channel(1):
        assert L._tp_channel[0] == 1
    assert L._tp_channel[0] == 0
    assert lib.ql_tp_set_concurrent(0) == 0


@pytest.mark.skipif(torch.cuda.is_available(), reason="CPU-box behaviour")
def test_no_cpu_fallback(lib):
    from nanovllm import _ext
    from nanovllm.utils import quantization as Q
    assert lib.ql_device_check() != 0
    x = torch.randn(4, 128).to(torch.bfloat16)
    with pytest.raises(RuntimeError):
        Q.per_row_quant(x, torch.int8)
    qw = torch.zeros(8, 128, dtype=torch.int8)
    with pytest.raises(RuntimeError):
        _ext.linear(x, qw, torch.ones(8), None, torch.int8, 0)
    with pytest.raises(AssertionError):                      # reference :151
        Q.per_row_quant(x, torch.float16)


# ------------------------------------------------------------------ TP host logic over gloo

def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _oracle_ext_standins():
    """CPU stand-ins for nanovllm._ext.{quant_rows, quant_groups, linear} built on oracle/."""
    from oracle import oracle as orc

    def quant_rows(x, dtype):
        q, s = orc.quant_rows(orc.t2n(x), orc.fmt_of(dtype))
        return orc.n2t(q, dtype), torch.from_numpy(s)

    def quant_groups(w, g, dtype):
        q, s = orc.quant_groups(orc.t2n(w), g, orc.fmt_of(dtype))
        return orc.n2t(q, dtype), torch.from_numpy(s)

    def linear(x, qw, sw, bias, dtype, group_size=0, out=None):
        fmt = orc.fmt_of(dtype)
        x2 = x.reshape(-1, x.shape[-1]).contiguous()
        qx, sx = orc.quant_rows(orc.t2n(x2), fmt)
        b = None if bias is None else orc.t2n(bias)
        if group_size:
            y = orc.gemm_groups(qx, sx, orc.t2n(qw), sw.numpy(), b, group_size, fmt)
        else:
            y = orc.gemm_rows(qx, sx, orc.t2n(qw), sw.numpy(), b, fmt)
        return orc.n2t(y, torch.bfloat16).view(list(x.shape[:-1]) + [qw.shape[0]])

    return quant_rows, quant_groups, linear


DIMS = dict(hidden=256, hd=64, heads=4, kv=2, inter=512)


def _full_weights():
    """Seeded full (unsharded) bf16 weights of one decoder layer's four linears."""
    d = DIMS
    g = torch.Generator().manual_seed(0)

    def rn(*shape, scale=0.05):
        return (torch.randn(*shape, generator=g, dtype=torch.float32) * scale).to(torch.bfloat16)

    return {"q": rn(d["heads"] * d["hd"], d["hidden"]), "k": rn(d["kv"] * d["hd"], d["hidden"]),
            "v": rn(d["kv"] * d["hd"], d["hidden"]), "o": rn(d["hidden"], d["heads"] * d["hd"]),
            "gate": rn(d["inter"], d["hidden"]), "up": rn(d["inter"], d["hidden"]),
            "down": rn(d["hidden"], d["inter"]), "o_bias": rn(d["hidden"], scale=1.0)}


def _tokens():
    return torch.randn(5, DIMS["hidden"], generator=torch.Generator().manual_seed(1),
                       dtype=torch.float32).to(torch.bfloat16)


def _tp_worker(rank, world, port, quant_type, ret):
    import torch.distributed as dist
    sys.path.insert(0, ROOT)
    sys.path.insert(0, os.path.join(ROOT, "nano-vllm-hw3_b200"))
    dist.init_process_group("gloo", init_method=f"tcp://127.0.0.1:{port}", rank=rank, world_size=world)
    try:
        from nanovllm import _ext
        _ext.quant_rows, _ext.quant_groups, _ext.linear = _oracle_ext_standins()
        from nanovllm.layers.linear import (MergedColumnParallelLinear, QKVParallelLinear,
                                            RowParallelLinear)
        torch.set_default_dtype(torch.bfloat16)
        hidden, hd, heads, kv, inter = (DIMS[k] for k in ('hidden', 'hd', 'heads', 'kv', 'inter'))
        full = _full_weights()
        qkv = QKVParallelLinear(hidden, hd, heads, kv)
        o = RowParallelLinear(heads * hd, hidden, bias=True)
        gate_up = MergedColumnParallelLinear(hidden, [inter, inter])
        down = RowParallelLinear(inter, hidden)
        assert (qkv.tp_rank, qkv.tp_size) == (rank, world)
        for sid in "qkv":
            qkv.weight.weight_loader(qkv.weight, full[sid], sid)
        o.weight.weight_loader(o.weight, full["o"])
        o.bias.weight_loader(o.bias, full["o_bias"])
        gate_up.weight.weight_loader(gate_up.weight, full["gate"], 0)
        gate_up.weight.weight_loader(gate_up.weight, full["up"], 1)
        down.weight.weight_loader(down.weight, full["down"])
        shards = dict(qkv=qkv.weight.data.clone(), o=o.weight.data.clone(),
                      gate_up=gate_up.weight.data.clone(), down=down.weight.data.clone())
        for m in (qkv, o, gate_up, down):
            m.quantize_(torch.int8, quant_type, 128)
            assert m.weight is None and m.qw.dtype == torch.int8
        x = _tokens()
        h = qkv(x)
        y_o = o(h[:, : heads * hd // world].contiguous())
        h2 = gate_up(x)
        y_d = down(h2[:, : inter // world].contiguous())
        ret[rank] = dict(shards=shards, h=h, y_o=y_o, h2=h2, y_d=y_d,
                         sw_shapes={n: tuple(m.sw.shape) for n, m in
                                    (("qkv", qkv), ("o", o), ("gate_up", gate_up), ("down", down))})
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("quant_type", ["per_row", "per_group"])
def test_tp2_gloo_sharding_bias_and_allreduce(quant_type):
    """world_size 2: loaders slice like linear.py:65-70,87-93,114-128,142-147; the quantized
    row-parallel linear adds its bias on rank 0 only and all-reduces (linear.py:150-152)."""
    import torch.multiprocessing as mp
    world, port = 2, _free_port()
    mgr = mp.Manager()
    ret = mgr.dict()
    mp.spawn(_tp_worker, args=(world, port, quant_type, ret), nprocs=world, join=True)
    r0, r1 = ret[0], ret[1]
    hidden, hd, heads, kv, inter = (DIMS[k] for k in ('hidden', 'hd', 'heads', 'kv', 'inter'))
    full = _full_weights()
    q, k, v, o_w, o_bias = full['q'], full['k'], full['v'], full['o'], full['o_bias']
    # column-parallel shards: rank r holds [q_r; k_r; v_r]
    for r, res in ((0, r0), (1, r1)):
        exp = torch.cat([q.chunk(2, 0)[r], k.chunk(2, 0)[r], v.chunk(2, 0)[r]], 0)
        assert torch.equal(res["shards"]["qkv"], exp)
        assert torch.equal(res["shards"]["o"], o_w.chunk(2, 1)[r])
        assert res["shards"]["gate_up"].shape == (2 * inter // 2, hidden)
        assert res["shards"]["down"].shape == (hidden, inter // 2)
    # scale layouts per shard
    if quant_type == "per_row":
        assert r0["sw_shapes"] == dict(qkv=(256,), o=(256,), gate_up=(512,), down=(256,))
    else:
        assert r0["sw_shapes"] == dict(qkv=(256, 2), o=(256, 1), gate_up=(512, 2), down=(256, 2))
    # all-reduced outputs are identical on both ranks
    assert torch.equal(r0["y_o"], r1["y_o"]) and torch.equal(r0["y_d"], r1["y_d"])
    # and equal the sum of the per-rank partials recomputed here (bias once, from rank 0)
    quant_rows, quant_groups, linear = _oracle_ext_standins()
    parts = []
    for r, res in ((0, r0), (1, r1)):
        w = res["shards"]["o"]
        if quant_type == "per_row":
            qw, sw = quant_rows(w, torch.int8)
            gs = 0
        else:
            qw, sw = quant_groups(w, 128, torch.int8)
            gs = 128
        xin = res["h"][:, : heads * hd // 2].contiguous()
        parts.append((xin, qw, sw, gs))
    y0 = linear(parts[0][0], parts[0][1], parts[0][2], o_bias, torch.int8, parts[0][3])
    y1 = linear(parts[1][0], parts[1][1], parts[1][2], None, torch.int8, parts[1][3])
    assert torch.equal(r0["y_o"], y0 + y1)
    # sanity: close to the unquantized full-precision result
    xq = _tokens()
    assert r0["h"].shape == (5, (heads + 2 * kv) * hd // 2)
    ref_q = xq.float() @ q.chunk(2, 0)[0].float().T
    got_q = r0["h"][:, : heads * hd // 2].float()
    assert (got_q - ref_q).abs().max() / ref_q.abs().max() < 0.08   # reference test threshold :549-553
