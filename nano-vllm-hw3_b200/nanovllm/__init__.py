"""B200-native quantized-linear engine behind nano-vllm's layer / quantization API.

Only the hot path lives here: ``nanovllm.utils.quantization`` and
``nanovllm.layers.linear`` mirror the reference modules of the same names; everything they
compute on a GPU goes through ``nanovllm._ext`` -> libqlinear_b200.so (sm_100a CUDA).
"""
__all__ = ["_ext"]
