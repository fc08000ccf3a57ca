"""Model-level parity (SURVEY.md section 8c level L4) and the plugin boundary, executed the way
the reference engine does it: ``apply_quantization`` (engine/model_runner.py:36-57) ->
``apply_per_row_quant / apply_group_quant / apply_tensor_quant`` (utils/quantization.py:463-496)
-> the in-place TP layers or, for a foreign LinearBase, ``QLinear / QGroupLinear /
QTensorLinear.from_linear`` (:349-401) -- on a test-only 4-layer Qwen3 (tests/tiny_qwen3.py,
Qwen3-0.6B widths, seeded random init) over a seeded synthetic token batch.

The comparison model runs the SAME non-linear GPU ops and takes its four projections from the CPU
oracle (``OracleLinear``), so every difference in the logits is the quantized-linear path's.

Bars (north_star: "bit-exact int32 accumulators for INT8 ... unchanged test_ppl.py perplexity
within 0.5%"; no Qwen3 weights / WikiText-2 exist offline, so the corpus is drawn from the bf16
teacher's own next-token distribution and the perplexity is taken in expectation over that
corpus -- tiny_qwen3.pseudo_ppl; the single-sample value is recorded next to it):

  INT8 per-row ............ logits BIT-EQUAL to the restated reference chain (every dispatch
                            branch that lands on it, TP-in-place and foreign-wrapped, fused
                            producers on and off, prefill- and decode-sized batches)
  FP8 per-row ............. pseudo-PPL within PPL_TOL = 0.5 % of the restated reference chain
  per-group / per-tensor .. pseudo-PPL within 0.5 % of the oracle's W8A8 chain (same semantics);
                            against the reference's weight-only semantics (activations kept
                            fp32, per_group_matmul :265-307) the W8A8 design differs by
                            construction (BASELINE.json configs[2] asks for per-token activation
                            quantisation): measured and bounded by W8A8_VS_WEIGHT_ONLY_TOL
  fused producers ......... logits bit-equal to the unfused path for every format

Measured numbers are written to gpurun_out/model_parity.json (copied into DESIGN.md section 5).
"""
import json
import os

import pytest
import torch

from tests import tiny_qwen3 as tq

pytestmark = pytest.mark.gpu

PPL_TOL = 0.005                   # north_star: within 0.5 %
W8A8_VS_WEIGHT_ONLY_TOL = 0.06    # stated bound for the deliberate W8A8-vs-weight-only deviation
LAYERS, VOCAB, BATCH, SEQ = 4, 2048, 4, 64
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
_RESULTS = {}


@pytest.fixture(scope="module")
def ext():
    assert torch.cuda.is_available(), "gpu tests need a CUDA device"
    from nanovllm import _ext
    assert _ext.load().ql_device_check() == 0, "device is not sm_100"
    return _ext


@pytest.fixture(scope="module", autouse=True)
def _dump_results():
    yield
    if _RESULTS:
        out = os.path.join(ROOT, "gpurun_out")
        try:
            os.makedirs(out, exist_ok=True)
            with open(os.path.join(out, "model_parity.json"), "w") as f:
                json.dump(_RESULTS, f, indent=1, sort_keys=True)
        except OSError:
            pass


def _ids(batch=BATCH, seq=SEQ):
    g = torch.Generator().manual_seed(7)
    return torch.randint(0, VOCAB, (batch, seq), generator=g).cuda()


@pytest.fixture(scope="module")
def corpus(ext):
    """bf16 teacher logits + the synthetic targets drawn from them."""
    teacher = tq.TinyQwen3ForCausalLM(LAYERS, VOCAB, tq.tp_linears)
    logits = teacher(_ids())
    targets = tq.synthetic_corpus(logits)
    return {"teacher_logits": logits.cpu(), "targets": targets, "ppl": tq.pseudo_ppl(logits, logits),
            "ppl_sampled": tq.pseudo_ppl_sampled(logits, targets)}


def _engine(factory, dtype, quant_type, group_size=128):
    from nanovllm.utils.quantization import apply_quantization
    m = tq.TinyQwen3ForCausalLM(LAYERS, VOCAB, factory)
    launches0 = None
    apply_quantization(m, dtype, quant_type, group_size)
    return m, launches0


def _oracle(dtype, quant_type, semantics, group_size=128):
    m = tq.TinyQwen3ForCausalLM(LAYERS, VOCAB, tq.foreign_linears)
    qt = quant_type or "per_row"
    m.swap_linears(lambda w: tq.OracleLinear(w, dtype, qt, group_size, semantics))
    return m


def _rel_l2(a, b):
    return float((a.double() - b.double()).norm() / b.double().norm())


# ------------------------------------------------------------------ INT8 per-row: bit-equal logits

@pytest.mark.parametrize("factory", ["tp", "foreign"])
@pytest.mark.parametrize("quant_type", ["per_row", None])
def test_int8_per_row_logits_bit_equal(ext, corpus, factory, quant_type):
    """Both ``quant_type == "per_row"`` (model_runner.py:46-47) and the default rule
    ``linear_dtype != bf16`` alone (:55-57) must produce the restated reference chain bit for bit."""
    from nanovllm.layers.linear import LinearBase
    from nanovllm.utils.quantization import QLinear
    fac = tq.tp_linears if factory == "tp" else tq.foreign_linears
    n0 = ext.launch_count()
    m, _ = _engine(fac, torch.int8, quant_type)
    for layer in m.model.layers:
        for lin in (layer.self_attn.qkv_proj, layer.self_attn.o_proj, layer.mlp.gate_up_proj,
                    layer.mlp.down_proj):
            if factory == "tp":
                assert isinstance(lin, LinearBase) and lin.is_quantized and lin.weight is None
                assert lin.quant_type == "per_row" and lin.qw.dtype == torch.int8
            else:
                assert isinstance(lin, QLinear) and lin.qw.dtype == torch.int8 and lin.sw.dtype == torch.float32
    ids = _ids()
    logits = m(ids)
    assert ext.launch_count() > n0, "the CUDA path did not run"
    ref = _oracle(torch.int8, quant_type, "reference")(ids)
    assert torch.equal(logits, ref), f"max |d logit| = {(logits - ref).abs().max().item()}"
    ppl = tq.pseudo_ppl(logits, corpus["teacher_logits"])
    _RESULTS[f"int8/per_row/{factory}/{quant_type}"] = {
        "ppl": ppl, "ppl_reference_chain": tq.pseudo_ppl(ref, corpus["teacher_logits"]),
        "ppl_sampled": tq.pseudo_ppl_sampled(logits, corpus["targets"]),
        "ppl_bf16_teacher": corpus["ppl"], "logits_bit_equal": True}


@pytest.mark.parametrize("batch,seq", [(4, 4), (2, 1), (3, 40)])
def test_int8_per_row_decode_sized_batches_bit_equal(ext, batch, seq):
    """M = 16, 2, 120 tokens: the decode-regime kernels (cluster split-K) inside the model."""
    m, _ = _engine(tq.tp_linears, torch.int8, "per_row")
    ids = _ids(batch, seq)
    assert torch.equal(m(ids), _oracle(torch.int8, "per_row", "reference")(ids))


# ------------------------------------------------------------------ fused producers == unfused

@pytest.mark.parametrize("dtype", [torch.int8, torch.float8_e4m3fn])
@pytest.mark.parametrize("quant_type", ["per_row", "per_group", "per_tensor"])
def test_fused_producer_path_equals_unfused(ext, dtype, quant_type):
    """RMSNorm.forward_quant / SiluAndMul.forward_quant -> QuantizedActivation -> linear must give
    the logits of the unfused chain exactly (the fused kernels quantise their own bf16 result)."""
    m, _ = _engine(tq.tp_linears, dtype, quant_type)
    ids = _ids()
    assert torch.equal(m(ids, fused=True, qdtype=dtype), m(ids))


# ------------------------------------------------------------------ every other branch: pseudo-PPL

@pytest.mark.parametrize("factory", ["tp", "foreign"])
@pytest.mark.parametrize("dtype_name", ["int8", "fp8"])
@pytest.mark.parametrize("quant_type", ["per_row", "per_group", "per_tensor"])
def test_pseudo_ppl_within_bar(ext, corpus, factory, dtype_name, quant_type):
    from nanovllm.utils import quantization as Q
    if dtype_name == "int8" and quant_type == "per_row":
        pytest.skip("covered bit-exactly by test_int8_per_row_logits_bit_equal")
    dtype = torch.int8 if dtype_name == "int8" else torch.float8_e4m3fn
    fac = tq.tp_linears if factory == "tp" else tq.foreign_linears
    m, _ = _engine(fac, dtype, quant_type)
    if factory == "foreign":
        want = {"per_row": Q.QLinear, "per_group": Q.QGroupLinear, "per_tensor": Q.QTensorLinear}[quant_type]
        assert all(isinstance(layer.mlp.down_proj, want) for layer in m.model.layers)
    ids = _ids()
    logits = m(ids)
    tl, tg = corpus["teacher_logits"], corpus["targets"]
    ppl = tq.pseudo_ppl(logits, tl)
    same = _oracle(dtype, quant_type, "w8a8")(ids)            # same semantics, restated on the CPU
    ppl_same = tq.pseudo_ppl(same, tl)
    rec = {"ppl": ppl, "ppl_oracle_w8a8": ppl_same, "ppl_bf16_teacher": corpus["ppl"],
           "ppl_vs_oracle_w8a8": ppl / ppl_same - 1.0, "logits_rel_l2_vs_oracle_w8a8": _rel_l2(logits, same),
           "ppl_sampled": tq.pseudo_ppl_sampled(logits, tg),
           "ppl_sampled_oracle_w8a8": tq.pseudo_ppl_sampled(same, tg)}
    assert abs(ppl / ppl_same - 1.0) <= PPL_TOL, rec
    if quant_type != "per_row":
        ref = _oracle(dtype, quant_type, "reference")(ids)    # the reference's weight-only semantics
        ppl_ref = tq.pseudo_ppl(ref, tl)
        rec.update(ppl_reference_weight_only=ppl_ref, ppl_vs_reference_weight_only=ppl / ppl_ref - 1.0,
                   logits_rel_l2_vs_reference_weight_only=_rel_l2(logits, ref))
        assert abs(ppl / ppl_ref - 1.0) <= W8A8_VS_WEIGHT_ONLY_TOL, rec
    _RESULTS[f"{dtype_name}/{quant_type}/{factory}"] = rec


# ------------------------------------------------------------------ dispatch corners

def test_dispatch_bf16_is_untouched_and_unknown_type_raises(ext):
    from nanovllm.utils.quantization import apply_quantization
    m = tq.TinyQwen3ForCausalLM(2, VOCAB, tq.tp_linears)
    apply_quantization(m, torch.bfloat16, None, 128)          # model_runner.py:55: nothing to do
    assert all(not layer.mlp.down_proj.is_quantized for layer in m.model.layers)
    with pytest.raises(ValueError):
        apply_quantization(m, torch.int8, "per_banana", 128)
    with pytest.raises(ValueError):
        apply_quantization(m, torch.int8, "smoothquant", 128)  # torchao hook: out of scope, refused loudly
    assert all(not layer.mlp.down_proj.is_quantized for layer in m.model.layers)
    # lm_head / embedding are never touched by the appliers (quantization.py:463-496)
    apply_quantization(m, torch.float8_e4m3fn, "per_group", 256)
    assert m.lm_head.weight.dtype == torch.bfloat16
    d = m.model.layers[0].mlp.down_proj
    assert d.quant_type == "per_group" and d.group_size == 256 and d.sw.shape == (1024, 3072 // 256)


def test_foreign_tp_module_is_refused_like_the_reference(ext):
    """QLinear.from_linear asserts tp_size == 1 (quantization.py:360)."""
    from nanovllm.utils.quantization import QGroupLinear, QLinear, QTensorLinear
    lin = tq.ForeignLinear(256, 128).cuda().to(torch.bfloat16)
    lin.tp_size = 2
    for make in (lambda: QLinear.from_linear(lin, torch.int8),
                 lambda: QTensorLinear.from_linear(lin, torch.int8),
                 lambda: QGroupLinear.from_linear(lin, torch.int8, 128)):
        with pytest.raises(AssertionError):
            make()
