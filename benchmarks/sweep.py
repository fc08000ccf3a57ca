#!/usr/bin/env python
"""Shape sweep of the quantized linear (BASELINE.json configs[4]).

    python benchmarks/sweep.py [--models 0.6b,8b] [--M 1,16,128,1024,8192] [--modes i8_row,f8_row,f8_g128,i8_g128]
                               [--out gpurun_out/sweep.json] [--split]

Each point times ql_linear (C ABI, through nanovllm._ext.linear) on one [N,K] shape of a Qwen3
model at M tokens: a CUDA graph holding one call per distinct weight copy is replayed, with
enough copies (>= 256 MB of weights + activations) that nothing is served from L2 between
uses; CUDA events around `iters` replays.  Reports us/call, TFLOP/s, algorithmic GB/s and the
roofline fraction t_min / t with t_min = max(FLOPs / P_tensor, bytes / BW_hbm) (peaks from
MEASURED_PEAKS.json: HBM copy GB/s; tensor = 2 x bf16 burst).  --split also times the
quantiser and the GEMM separately for M > 16.
"""
import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for _p in (ROOT, os.path.join(ROOT, "nano-vllm-hw3_b200")):
    if _p not in sys.path:
        sys.path.insert(0, _p)

import torch  # noqa: E402

from bench import MODELS, layer_shapes, load_peaks  # noqa: E402

MODES = {"i8_row": (torch.int8, 0), "f8_row": (torch.float8_e4m3fn, 0),
         "i8_g128": (torch.int8, 128), "f8_g128": (torch.float8_e4m3fn, 128)}


def timed_graph(fn, iters=20, warm=3):
    s = torch.cuda.Stream()
    s.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(s):
        fn()
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g, stream=s):
            fn()
    torch.cuda.current_stream().wait_stream(s)
    for _ in range(warm):
        g.replay()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        g.replay()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / iters * 1e-3


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--models", default="0.6b,8b")
    ap.add_argument("--M", default="1,16,128,1024,8192")
    ap.add_argument("--modes", default="i8_row,f8_row,f8_g128,i8_g128")
    ap.add_argument("--out", default=os.path.join(ROOT, "gpurun_out", "sweep.json"))
    ap.add_argument("--split", action="store_true")
    ap.add_argument("--working-set-mb", type=int, default=256)
    args = ap.parse_args()
    from nanovllm import _ext
    peaks = load_peaks()
    P = 2.0 * peaks["bf16"] * 1e12
    BW = peaks["hbm"] * 1e9
    rows = []
    print(f"{'model':5s} {'lin':8s} {'mode':8s} {'M':>5s} {'N':>6s} {'K':>6s} {'us':>9s} {'TFLOP/s':>9s} "
          f"{'GB/s':>8s} {'frac':>6s} bound  path")
    for model in args.models.split(","):
        for name, N, K in layer_shapes(MODELS[model]):
            for mode in args.modes.split(","):
                dtype, g = MODES[mode]
                for M in [int(v) for v in args.M.split(",")]:
                    per_copy = N * K + 2 * M * K + 2 * M * N
                    copies = max(2, min(64, -(-args.working_set_mb * (1 << 20) // per_copy)))
                    gen = torch.Generator(device="cuda").manual_seed(N + K + M)
                    ws, xs, ys = [], [], []
                    for _ in range(copies):
                        w = (torch.randn(N, K, device="cuda", generator=gen) * 0.02).to(torch.bfloat16)
                        q, s = _ext.quant_groups(w, g, dtype) if g else _ext.quant_rows(w, dtype)
                        ws.append((q, s))
                        xs.append(torch.randn(M, K, device="cuda", generator=gen).to(torch.bfloat16))
                        ys.append(torch.empty(M, N, device="cuda", dtype=torch.bfloat16))
                        del w

                    def run():
                        for (q, s), x, y in zip(ws, xs, ys):
                            _ext.linear(x, q, s, None, dtype, g, out=y)

                    t = timed_graph(run) / copies
                    fl = 2.0 * M * N * K
                    sw = 4 * N * (K // g if g else 1)
                    by = (2 * M * K + N * K + sw + 2 * M * N) if M <= 16 else \
                         (3 * M * K + 4 * M) + (M * K + N * K + 4 * M + sw + 2 * M * N)
                    t_fl, t_by = fl / P, by / BW
                    rec = dict(model=model, linear=name, mode=mode, M=M, N=N, K=K, us=t * 1e6,
                               tflops=fl / t / 1e12, gbs=by / t / 1e9, frac=max(t_fl, t_by) / t,
                               bound="tensor" if t_fl >= t_by else "hbm", copies=copies,
                               path=_ext.linear_path(M, N, K, g))
                    if args.split:
                        qs = [_ext.quant_rows(x, dtype) for x in xs]

                        def run_q():
                            for x in xs:
                                _ext.quant_rows(x, dtype)

                        def run_g():
                            for (q, s), (qx, sx) in zip(ws, qs):
                                _ext.gemm_w8a8(qx, sx, q, s, None, dtype, g)

                        def run_qg():
                            for (q, s), x in zip(ws, xs):
                                qx, sx = _ext.quant_rows(x, dtype)
                                _ext.gemm_w8a8(qx, sx, q, s, None, dtype, g)

                        rec["chain_us"] = timed_graph(run_qg) / copies * 1e6
                        rec["quant_us"] = timed_graph(run_q) / copies * 1e6
                        rec["gemm_us"] = timed_graph(run_g) / copies * 1e6
                        rec["gemm_tflops"] = fl / (rec["gemm_us"] * 1e-6) / 1e12
                    rows.append(rec)
                    extra = (f"  q+g chained {rec['chain_us']:.1f}us quant {rec['quant_us']:.1f}us gemm {rec['gemm_us']:.1f}us "
                             f"({rec['gemm_tflops']:.0f} TF)") if "gemm_us" in rec else ""
                    print(f"{model:5s} {name:8s} {mode:8s} {M:5d} {N:6d} {K:6d} {rec['us']:9.2f} "
                          f"{rec['tflops']:9.1f} {rec['gbs']:8.0f} {rec['frac']:6.3f} {rec['bound']:6s} "
                          f"{rec['path']}{extra}", flush=True)
                    del ws, xs, ys
                    torch.cuda.empty_cache()
    os.makedirs(os.path.dirname(args.out), exist_ok=True)
    json.dump(dict(peaks=peaks, tensor_peak_tflops=P / 1e12, rows=rows), open(args.out, "w"), indent=1)


if __name__ == "__main__":
    main()
