#!/usr/bin/env python
"""bench.py -- the quantized-linear hot path of Qwen3 through the B200-native engine.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl b200|reference]
                    [--model 0.6b|1.7b|8b] [--tokens T] [--dtype int8|fp8] [--quant per_row|per_group]

A "step" is one pass of the hot path over one batch of synthetic tokens: all 4 x n_layers
quantized linears of the model (qkv -> o -> gate_up -> down per layer, distinct random-init
weights per layer, so one pass streams every weight once and the working set exceeds L2).
Default workload = BASELINE.json configs[1]/[0] shapes: Qwen3-0.6B, W8A8 INT8 per-row,
T = 4096 tokens (8 x 512) per rank.

N > 1 (launched by torchrun, one rank per GPU): the model is tensor-parallel TP = N exactly
as nano-vllm shards it (column-parallel qkv/gate_up, row-parallel o/down + NCCL all-reduce);
the batch grows with N (T x N tokens, "weak"), every rank sees every token.

One JSON line on rank 0 (contract in the task statement): value = tokens/s with inputs
resident in HBM; e2e = same metric with HOST (pinned) input/output buffers and the copies
inside the timed region; roofline = the dominant kernel (tcgen05 GEMM) timed alone with CUDA
events over the same weights; cpu_baseline = the reference's CPU path (torch dequant+matmul,
oracle.ref_per_group_matmul) on a bounded sample.  --impl reference prints the CPU arm only.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import tempfile
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
for _p in (ROOT, os.path.join(ROOT, "nano-vllm-hw3_b200")):
    if _p not in sys.path:
        sys.path.insert(0, _p)

import torch  # noqa: E402

MODELS = {  # hidden, intermediate, layers, heads, kv_heads, head_dim   
    "0.6b": dict(hidden=1024, inter=3072, layers=28, heads=16, kv=8, hd=128),
    "1.7b": dict(hidden=2048, inter=6144, layers=28, heads=16, kv=8, hd=128),
    "8b": dict(hidden=4096, inter=12288, layers=36, heads=32, kv=8, hd=128),
}
METRIC = "qwen3_w8a8_linear_stack_tokens_per_s"


def load_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return dict(hbm=d["hbm_gbs"], bf16=d["bf16_tflops"], bf16_sus=d["bf16_tflops_sustained"],
                    src="measured (MEASURED_PEAKS.json)")
    return dict(hbm=6650.0, bf16=1590.0, bf16_sus=1400.0, src="fallback (B200_PROFILING.md)")


def load_traffic(key):
    """DRAM bytes per launch of the dominant kernel from the committed `ncu --set full` capture of
    this same command (profiles/traffic.json, written by tools/ncu_traffic.py); None when the
    capture was taken on another workload."""
    p = os.path.join(ROOT, "profiles", "traffic.json")
    if not os.path.exists(p):
        return None
    d = json.load(open(p))
    return d.get("traffic_bytes_per_launch") if d.get("workload", key) == key else None


def layer_shapes(m, tp=1):
    """[(name, N, K)] per rank for one decoder layer."""
    qkv_n = (m["heads"] + 2 * m["kv"]) * m["hd"]
    return [("qkv", qkv_n // tp, m["hidden"]), ("o", m["hidden"], m["heads"] * m["hd"] // tp),
            ("gate_up", 2 * m["inter"] // tp, m["hidden"]), ("down", m["hidden"], m["inter"] // tp)]


def flops_bytes(M, shapes, group=0):
    """Algorithmic work of one layer: FLOPs = 2MNK; bytes = quant
    (2MK + MK + 4M) + GEMM (MK + NK + 4M + S_w + 2MN) for M > 16, fused (2MK + NK + S_w + 2MN) else."""
    fl = by = 0
    for _, N, K in shapes:
        fl += 2 * M * N * K
        sw = 4 * N * (K // group if group else 1)
        by += (2 * M * K + N * K + sw + 2 * M * N) if M <= 16 else \
              (3 * M * K + 4 * M) + (M * K + N * K + 4 * M + sw + 2 * M * N)
    return fl, by


# ------------------------------------------------------------------ clocks sampler

class ClockSampler:
    """SM clock + throttle reasons sampled DURING the timed regions.  NVML in a background thread
    (2 ms period: the default run times only tens of milliseconds, far below nvidia-smi's start-up
    time); falls back to the profiling recipe's `nvidia-smi -lms` process when NVML is unusable."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,"
         "clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, dev_index: int):
        import threading
        self.samples = []           # (sm_mhz, reasons bitmask, power_w)
        self.sm_max = None
        self.mode = None
        self._stop = threading.Event()
        self.p = self.f = self.t = None
        try:
            import pynvml
            pynvml.nvmlInit()
            vis = os.environ.get("CUDA_VISIBLE_DEVICES")
            phys = dev_index
            if vis:
                tok = vis.split(",")[dev_index].strip()
                phys = int(tok) if tok.isdigit() else None
                h = pynvml.nvmlDeviceGetHandleByUUID(tok) if phys is None else pynvml.nvmlDeviceGetHandleByIndex(phys)
            else:
                h = pynvml.nvmlDeviceGetHandleByIndex(phys)
            self.sm_max = float(pynvml.nvmlDeviceGetMaxClockInfo(h, pynvml.NVML_CLOCK_SM))
            self._nv, self._h = pynvml, h
            self.mode = "nvml"
            self.t = threading.Thread(target=self._run, daemon=True)
            self.t.start()
        except Exception:
            self.mode = None
        if self.mode is None:
            try:
                self.f = tempfile.NamedTemporaryFile("w+", suffix=".csv", delete=False)
                self.p = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                           "-lms", "100", "-i", str(dev_index)], stdout=self.f,
                                          stderr=subprocess.DEVNULL)
                self.mode = "nvidia-smi"
            except OSError:
                self.p = None

    def _run(self):
        nv, h = self._nv, self._h
        while not self._stop.is_set():
            try:
                sm = nv.nvmlDeviceGetClockInfo(h, nv.NVML_CLOCK_SM)
                rs = nv.nvmlDeviceGetCurrentClocksEventReasons(h) if hasattr(
                    nv, "nvmlDeviceGetCurrentClocksEventReasons") else nv.nvmlDeviceGetCurrentClocksThrottleReasons(h)
                try:
                    pw = nv.nvmlDeviceGetPowerUsage(h) / 1000.0
                except Exception:
                    pw = 0.0
                self.samples.append((float(sm), int(rs), pw))
            except Exception:
                pass
            time.sleep(0.002)

    def stop(self):
        out = {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0, "sampler": self.mode}
        if self.mode == "nvml":
            self._stop.set()
            self.t.join(timeout=2)
            nv = self._nv
            bits = {"hw_slowdown": 0x8, "hw_thermal_slowdown": 0x40, "sw_thermal_slowdown": 0x20,
                    "sw_power_cap": 0x4}
            if self.samples:
                s = sorted(v[0] for v in self.samples)
                mask = 0
                for v in self.samples:
                    mask |= v[1]
                out.update(sm_mhz=s[len(s) // 2], sm_max_mhz=self.sm_max,
                           reasons=sorted(n for n, b in bits.items() if mask & b),
                           samples=len(s), power_w_max=max(v[2] for v in self.samples))
            return out
        if self.p is None:
            return out
        self.p.terminate()
        try:
            self.p.wait(timeout=5)
        except subprocess.TimeoutExpired:
            self.p.kill()
        self.f.flush()
        self.f.seek(0)
        sm, mx, reasons, pw = [], [], set(), []
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for line in self.f.read().splitlines():
            c = [t.strip() for t in line.split(",")]
            if len(c) < 9:
                continue
            try:
                sm.append(float(c[1])); mx.append(float(c[2])); pw.append(float(c[3]))
            except ValueError:
                continue
            for n, v in zip(names, c[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        os.unlink(self.f.name)
        if sm:
            s = sorted(sm)
            out.update(sm_mhz=s[len(s) // 2], sm_max_mhz=max(mx), reasons=sorted(reasons),
                       samples=len(sm), power_w_max=max(pw))
        return out


# ------------------------------------------------------------------ the stack under test

class LinearStack:
    """n_layers x (qkv, o, gate_up, down) built from the engine's own TP layer classes."""

    def __init__(self, m, dtype, quant, group, device, seed=0):
        from nanovllm.layers.linear import (MergedColumnParallelLinear, QKVParallelLinear,
                                            RowParallelLinear)
        self.m = m
        prev = torch.get_default_dtype()
        torch.set_default_dtype(torch.bfloat16)
        g = torch.Generator(device=device).manual_seed(seed)
        self.layers = []
        with torch.device(device):
            for _ in range(m["layers"]):
                mods = [QKVParallelLinear(m["hidden"], m["hd"], m["heads"], m["kv"]),
                        RowParallelLinear(m["heads"] * m["hd"], m["hidden"]),
                        MergedColumnParallelLinear(m["hidden"], [m["inter"]] * 2),
                        RowParallelLinear(m["inter"], m["hidden"])]
                for mod in mods:
                    mod.weight.data.normal_(0.0, 0.02, generator=g)   # random-init (no checkpoints offline)
                    mod.quantize_(dtype, quant, group)
                self.layers.append(mods)
        torch.set_default_dtype(prev)
        self.tp = self.layers[0][0].tp_size
        self.side = None

    @torch.inference_mode()
    def forward(self, x):
        """x bf16 [T, hidden] -> bf16 [T, hidden].  Attention / SiLU / norms are outside the
        hot path: the o_proj and down_proj inputs are the leading columns of the previous
        output (a strided view, exercising ldx); every layer consumes the previous layer's output,
        as the decoder stack does (qwen3.py:186-187).
        Under tensor parallelism a row-parallel output that feeds the next quantized linear takes
        the sequence-parallel hand-off (RowParallelLinear.forward_quant: GEMM-epilogue
        reduce-scatter -> quantise the owner's tokens -> all-gather codes + scales); the last
        linear returns the all-reduced bf16 tensor (RowParallelLinear.forward)."""
        T = x.shape[0]
        nmb = self.micro_batches(T)
        if nmb == 1:
            return self._pass(x)
        # Two token micro-batches on two streams, one collective channel each: the NVLink time,
        # flag latencies and write acknowledgements of one micro-batch's reduce-scatter / all-gather
        # hide under the tensor-core work of the other (include/qlinear_b200.h, "PER CHANNEL").
        from nanovllm.layers.linear import tp_channel
        cur = torch.cuda.current_stream()
        if self.side is None:
            self.side = [torch.cuda.Stream() for _ in range(nmb)]
        cut = (T // nmb + 255) // 256 * 256                 # whole 256-token tiles per micro-batch
        outs = []
        for c in range(nmb):
            st = self.side[c]
            st.wait_stream(cur)
            with torch.cuda.stream(st), tp_channel(c):
                outs.append(self._pass(x[c * cut: min(T, (c + 1) * cut)]))
        for st in self.side:
            cur.wait_stream(st)
        return torch.cat(outs)

    def micro_batches(self, T):
        """QL_TP_MICRO=2: two token micro-batches on two streams (one NVLink collective channel each)
        when the step is tensor-parallel, prefill-sized and the NVLink path with >= 2 channels is up.
        Default 1.  Measured (profiles/r2e_*, r2j_*): the split loses 6-9 % at TP = 2 / 4 on
        Qwen3-0.6B (half-size GEMMs cost more than the overlap wins) and won 6 % in one TP = 8 run
        -- but a second TP = 8 run of the same configuration ended in the bounded-wait trap (a peer
        flag never arrived: an ordering hazard between the two channels that the single-channel
        path does not have), so it stays opt-in."""
        want = os.environ.get("QL_TP_MICRO", "1")
        if self.tp == 1 or want != "2" or self.layers[0][1].group_size:
            return 1
        if os.environ.get("QL_TP_ALLOW_CONCURRENT", "")[:1] != "1":
            return 1     # the library refuses two channels without this second, explicit opt-in
        from nanovllm import _comm
        comm = _comm.get()
        if comm is None or comm.channels < 2 or T < 2 * 256 * self.tp:
            return 1
        half = (T // 2 + 255) // 256 * 256
        hid = self.m["hidden"]
        return 2 if (comm.fits_big(half, hid) and not comm.fits(half * hid)) else 1

    def _pass(self, x):
        m = self.m
        k_o, k_down = m["heads"] * m["hd"] // self.tp, m["inter"] // self.tp
        h_in, last = x, len(self.layers) - 1
        for i, (qkv, o, gate_up, down) in enumerate(self.layers):
            h = qkv(h_in)
            h = gate_up(o.forward_quant(h[:, :k_o]) if self.tp > 1 else o(h[:, :k_o]))
            h = h[:, :k_down]
            h_in = down.forward_quant(h) if (self.tp > 1 and i != last) else down(h)
        return h_in


def time_steps(fn, steps, warmup, dist_on, finish=None):
    import torch.distributed as dist
    for _ in range(warmup):
        fn()
    if finish is not None:
        finish()
    if dist_on:
        dist.barrier()
    torch.cuda.synchronize()
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(2)]
    ev[0].record()
    for _ in range(steps):
        fn()
    if finish is not None:
        finish()                                           # the last step's D2H is inside the timed region
    ev[1].record()
    torch.cuda.synchronize()
    ms = ev[0].elapsed_time(ev[1])
    if dist_on:
        t = torch.tensor([ms], device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        dist.barrier()
        ms = float(t.item())
    return ms / steps


# ------------------------------------------------------------------ TP check, decode leg

def tp_check(stack, T, dtype, rank, world):
    """Run the collectives of the timed path on layer 0's o_proj and compare with their definition.
    Raises (non-zero exit) on any mismatch; returns what ran."""
    import torch.distributed as dist
    from nanovllm import _comm, _ext
    o = stack.layers[0][1]
    comm = _comm.get()
    info = {"status": "ok", "world": world, "checked": []}
    K = o.qw.shape[1]
    N = o.qw.shape[0]
    for M in (T, 16 * world):
        g = torch.Generator(device="cuda").manual_seed(100 + rank)
        x = torch.randn(M, K, generator=g, device="cuda", dtype=torch.float32).to(torch.bfloat16)
        part = o.forward_partial(x)
        parts = [torch.empty_like(part) for _ in range(world)]
        dist.all_gather(parts, part)
        acc = parts[0].float()
        for p in parts[1:]:
            acc = acc + p.float()
        y_def = acc.to(torch.bfloat16)
        q_def, s_def = _ext.quant_rows(y_def, dtype)
        qa = o.forward_quant(x)
        y = o(x)
        torch.cuda.synchronize()
        if comm is None or not (comm.fits(M * N) or comm.fits_big(M, N)):
            which = "nccl all_reduce (bf16 ring: tolerance check)"
            ok = bool((y.float() - y_def.float()).abs().max() <= 2.0 ** -6 * y_def.float().abs().max())
        else:
            which = ("one-shot NVLink push all-reduce (ar_oneshot_kernel)" if comm.fits(M * N) else
                     ("gemm2_push_kernel (GEMM epilogue -> peer TMA stores) + tp_reduce_kernel" if
                      (comm.push and o.group_size == 0) else "tp_reduce_kernel (owners pull)"))
            ok = bool(torch.equal(y, y_def) and torch.equal(qa.s, s_def)
                      and torch.equal(qa.q.view(torch.uint8), q_def.view(torch.uint8)))
        info["checked"].append({"tokens": M, "hidden": N, "collective": which, "bit_identical": ok})
        if ok and comm is not None and comm.fits_fused(M, N):
            # the fused all-reduce + residual add + RMSNorm + quantiser (ar_rmsnorm_quant_kernel,
            # RMSNorm.forward_quant(partial=True), layernorm.py:28-40 behind linear.py:150-152): its
            # definition is ql_rmsnorm_quant on the rank-order-summed bf16 tensor
            from nanovllm.layers.layernorm import RMSNorm
            norm = RMSNorm(N).cuda()
            gn = torch.Generator(device="cuda").manual_seed(7)                      # same on every rank
            norm.weight.data = (1 + 0.1 * torch.randn(N, generator=gn, device="cuda")).to(torch.bfloat16)
            res = torch.randn(M, N, generator=gn, device="cuda").to(torch.bfloat16)
            qa2, res2 = norm.forward_quant(part, res, dtype, partial=True)
            qd, sd, _, resd = norm._run(y_def, res, dtype, False)
            torch.cuda.synchronize()
            ok = bool(torch.equal(res2, resd) and torch.equal(qa2.s, sd)
                      and torch.equal(qa2.q.view(torch.uint8), qd.view(torch.uint8)))
            info["checked"].append({"tokens": M, "hidden": N, "bit_identical": ok,
                                    "collective": "all-reduce + add + RMSNorm + quant in one kernel (ar_rmsnorm_quant_kernel)"})
        flag = torch.tensor([1 if ok else 0], device="cuda")
        dist.all_reduce(flag, op=dist.ReduceOp.MIN)
        if not bool(flag.item()):
            info["status"] = "MISMATCH"
            if rank == 0:
                print(json.dumps({"tp_check": "MISMATCH", "detail": info}), flush=True)
            torch.cuda.synchronize()
            os._exit(3)
    return info


def tp_8b_leg(dtype, quant, T_rank, rank, world, peaks, args):
    """BASELINE.json configs[3]: the Qwen3-8B linear stack, INT8 per-row, tensor
    parallel over all `world` ranks -- per-rank shards qkv [6144/tp, 4096], o [4096, 4096/tp], gate_up
    [24576/tp, 4096], down [4096, 12288/tp]; 72 row-parallel exchanges per pass.  Prefill at T_rank x
    world tokens and decode at 16 x world tokens, CUDA-graph replay, device-timed, max over ranks;
    beside the step time the local GEMMs alone (no quantiser, no collective), so the part of the step
    that is NOT tensor-core work is visible.  Random-init weights, synthetic activations."""
    from nanovllm import _ext
    m = MODELS["8b"]
    stack = LinearStack(m, dtype, quant, 0, "cuda", seed=3)
    shapes = layer_shapes(m, world)
    out = {"model": "qwen3-8b (36 layers, hidden 4096, intermediate 12288), INT8 per-row, random-init",
           "tp": world, "shards_per_rank": {n: [N, K] for n, N, K in shapes}}
    steps, warm = max(args.steps // 4, 3), 3

    def graph_of(fn):
        st = torch.cuda.Stream()
        st.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(st):
            fn()
            gr = torch.cuda.CUDAGraph()
            with torch.cuda.graph(gr, stream=st):
                fn()
        torch.cuda.current_stream().wait_stream(st)
        return gr

    for name, T in (("prefill", T_rank * world), ("decode", 16 * world)):
        g = torch.Generator(device="cuda").manual_seed(5)
        x = torch.randn(T, m["hidden"], generator=g, device="cuda", dtype=torch.float32).to(torch.bfloat16)
        n0 = _ext.launch_count()
        stack.forward(x)
        launches = _ext.launch_count() - n0
        fn = (lambda: stack.forward(x)) if args.no_graph else graph_of(lambda: stack.forward(x)).replay
        ms = time_steps(fn, steps, warm, True)
        acts = {}
        for (_, N, K) in shapes:
            if K not in acts:
                xa = torch.randn(T, K, device="cuda", dtype=torch.float32, generator=g).to(torch.bfloat16)
                acts[K] = _ext.quant_rows(xa, dtype)

        def gemms():
            for mods in stack.layers:
                for mod in mods:
                    q, sc = acts[mod.qw.shape[1]]
                    _ext.gemm_w8a8(q, sc, mod.qw, mod._sw_kernel, None, dtype, 0)

        ms_g = time_steps(gemms if args.no_graph else graph_of(gemms).replay, steps, warm, True)
        fl = sum(2 * T * N * K for _, N, K in shapes) * m["layers"]                 # per rank
        by = sum(N * K for _, N, K in shapes) * m["layers"]                         # weight bytes per rank
        leg = {"tokens_per_step": T, "ms_per_step": ms, "value": T / (ms * 1e-3), "unit": "tokens/s",
               "gpu_launches_per_step": int(launches), "local_gemms_ms": ms_g,
               "gemm_share_of_step": ms_g / ms, "not_gemm_ms": ms - ms_g,
               "exchanges_per_step": 2 * m["layers"],
               "bytes_per_exchange": T * m["hidden"] * 2}
        if name == "prefill":
            leg.update(tflops_per_rank=fl / (ms * 1e-3) / 1e12,
                       frac_of_tensor_peak=fl / (ms * 1e-3) / 1e12 / (2.0 * peaks["bf16"]),
                       frac_of_nominal_4500=fl / (ms * 1e-3) / 1e12 / 4500.0)
        else:
            leg.update(weight_gbs_per_rank=by / (ms * 1e-3) / 1e9,
                       frac_of_hbm_peak=by / (ms * 1e-3) / 1e9 / peaks["hbm"])
        out[name] = leg
    info = tp_check(stack, T_rank * world, dtype, rank, world)
    out["tp_check"] = info["status"]
    out["tp_collective"] = info["checked"]
    del stack
    torch.cuda.empty_cache()
    return out


def decode_leg(stack, m, T, dtype, group, world, peaks, args, dist_on):
    """The same stack at a decode-sized batch: HBM-bound (every weight byte is streamed once per
    step), reported against the measured HBM peak.  Single GPU, per-row scales: the whole stack is
    ONE persistent launch (ql_chain_run, csrc/decode_chain.cu); the per-linear launches
    (CUDA-graph replay of ql_linear x 4 x layers) are timed beside it and must give the same
    bits for INT8.  TP > 1 / per-group: per-linear launches only."""
    from nanovllm import _ext
    gx = torch.Generator(device="cuda").manual_seed(2)
    x = torch.randn(T, m["hidden"], generator=gx, device="cuda", dtype=torch.float32).to(torch.bfloat16)
    n0 = _ext.launch_count()
    y_ref = stack.forward(x)
    launches = _ext.launch_count() - n0
    if not args.no_graph:
        s = torch.cuda.Stream()
        s.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(s):
            gr = torch.cuda.CUDAGraph()
            with torch.cuda.graph(gr, stream=s):
                stack.forward(x)
        torch.cuda.current_stream().wait_stream(s)
        fn = gr.replay
    else:
        def fn():
            stack.forward(x)
    steps, warm = max(args.steps, 10), max(args.warmup, 3)
    ms_lin = time_steps(fn, steps, warm, dist_on)
    shapes = layer_shapes(m, world)
    _, by_layer = flops_bytes(T, shapes, group)
    by = by_layer * m["layers"]
    out = {"tokens_per_step": T, "bound": "hbm", "algorithmic_bytes_per_step": by, "peak_gbs": peaks["hbm"],
           "unit": "tokens/s",
           "per_linear_launches": {"ms_per_step": ms_lin, "gpu_launches_per_step": int(launches),
                                   "frac": by / (ms_lin * 1e-3) / 1e9 / peaks["hbm"],
                                   "kernel": "skinny_kernel (tcgen05 cluster split-K), one launch per linear, "
                                             "cuda graph replay"}}
    ms, kern, n_launch = ms_lin, out["per_linear_launches"]["kernel"], int(launches)
    if world == 1 and group == 0 and T <= 16 and os.environ.get("QL_NO_CHAIN", "0") != "1":
        ops = []
        for li, mods in enumerate(stack.layers):
            for j, mod in enumerate(mods):
                ops.append((mod.qw, mod._sw_kernel, mod.bias, x if (li == 0 and j == 0) else None))
        chain = _ext.DecodeChain(ops, T, dtype)
        y_chain = chain.run()
        torch.cuda.synchronize()
        same = bool(torch.equal(y_chain, y_ref))
        if dtype == torch.int8:
            assert same, "decode chain result differs from the per-linear path (INT8 must be bit-identical)"
        ms_chain = time_steps(chain.run, steps, warm, dist_on)
        out["chain"] = {"ms_per_step": ms_chain, "gpu_launches_per_step": 1,
                        "frac": by / (ms_chain * 1e-3) / 1e9 / peaks["hbm"],
                        "equals_per_linear_bits": same,
                        "kernel": "decode_chain_kernel (persistent, one launch per step)"}
        if ms_chain < ms:
            ms, kern, n_launch = ms_chain, out["chain"]["kernel"], 1
    out.update(ms_per_step=ms, value=T / (ms * 1e-3), gpu_launches_per_step=n_launch, kernel=kern,
               achieved_gbs=by / (ms * 1e-3) / 1e9, frac=by / (ms * 1e-3) / 1e9 / peaks["hbm"],
               note="whole step incl. activation quantisation, per rank; weights stream from HBM "
                    "(the stack's distinct weights exceed L2); headline = the faster of the two launch forms")
    return out


def model_shaped_leg(stack, m, T, dtype, args):
    """The decoder layer's data flow around the same 112 quantized linears, with the producers the
    model really has (reference models/qwen3.py:147-167, 114-119): add + RMSNorm -> qkv -> [attention
    stand-in: the leading k_o columns] -> o -> add + RMSNorm -> gate_up -> SiLU * mul -> down.
    `fused` = the producers hand the linears their input already quantised (ql_rmsnorm_quant,
    ql_silu_mul_quant: SURVEY 8f-1; 5 launches fewer per layer); `unfused` = reference structure
    (bf16 norm / activation, then the linear's own per_row_quant).  Same bits either way."""
    from nanovllm import _ext
    from nanovllm.layers.activation import SiluAndMul
    from nanovllm.layers.layernorm import RMSNorm
    prev = torch.get_default_dtype()
    torch.set_default_dtype(torch.bfloat16)
    with torch.device("cuda"):
        norms = [(RMSNorm(m["hidden"]), RMSNorm(m["hidden"])) for _ in stack.layers]
    torch.set_default_dtype(prev)
    g = torch.Generator(device="cuda").manual_seed(7)
    for a, b in norms:
        for n in (a, b):
            n.weight.data.copy_(1.0 + 0.1 * torch.randn(m["hidden"], generator=g, device="cuda"))
    act = SiluAndMul()
    k_o = m["heads"] * m["hd"]
    x = torch.randn(T, m["hidden"], generator=g, device="cuda", dtype=torch.float32).to(torch.bfloat16)

    def forward(fused):
        h, res = x, None
        for (qkv, o, gate_up, down), (ln1, ln2) in zip(stack.layers, norms):
            if fused:
                if res is None:
                    a, res = ln1.forward_quant(h, None, dtype), h
                else:
                    a, res = ln1.forward_quant(h, res, dtype)
            else:
                if res is None:
                    a, res = ln1(h), h
                else:
                    a, res = ln1(h, res)
            h = o(qkv(a)[:, :k_o])
            if fused:
                a, res = ln2.forward_quant(h, res, dtype)
                h = down(act.forward_quant(gate_up(a), dtype))
            else:
                a, res = ln2(h, res)
                h = down(act(gate_up(a)))
        return h

    out = {"tokens_per_step": T}
    ys = {}
    for name, fused in (("fused_producers", True), ("unfused", False)):
        n0 = _ext.launch_count()
        ys[name] = forward(fused)
        launches = _ext.launch_count() - n0
        s = torch.cuda.Stream()
        s.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(s):
            gr = torch.cuda.CUDAGraph()
            with torch.cuda.graph(gr, stream=s):
                forward(fused)
        torch.cuda.current_stream().wait_stream(s)
        ms = time_steps(gr.replay, max(args.steps // 2, 5), 3, False)
        out[name] = {"ms_per_step": ms, "value": T / (ms * 1e-3), "unit": "tokens/s",
                     "gpu_launches_per_step": int(launches)}
    out["same_bits"] = bool(torch.equal(ys["fused_producers"], ys["unfused"]))
    assert out["same_bits"], "fused producers changed the result"
    out["speedup"] = out["unfused"]["ms_per_step"] / out["fused_producers"]["ms_per_step"]
    return out


def fp8_group_leg(m, T, peaks, args):
    """BASELINE.json configs[2] / config 3 of SURVEY 8d: FP8-e4m3 weights with one scale per 128
    values of K (reference per_group_quant / per_group_matmul, quantization.py:194-234, 265-307), the
    same stack and token count, single GPU: prefill pass and a 16-token decode pass, CUDA-graph
    replay.  Per-group is bounded by the TMEM READ path, not the tensor pipe (every 128-byte K block
    the whole accumulator goes through tcgen05.ld to be promoted with its scale: 64 tensor FLOP per
    TMEM byte), so its fraction of the tensor roofline is expected well below the per-row one."""
    from nanovllm import _ext
    stack = LinearStack(m, torch.float8_e4m3fn, "per_group", 128, "cuda", seed=3)
    out = {"quant": "fp8_e4m3 per_group g=128"}
    gx = torch.Generator(device="cuda").manual_seed(4)
    for name, tok in (("prefill", T), ("decode", 16)):
        x = torch.randn(tok, m["hidden"], generator=gx, device="cuda", dtype=torch.float32).to(torch.bfloat16)
        s = torch.cuda.Stream()
        s.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(s):
            stack.forward(x)
            gr = torch.cuda.CUDAGraph()
            with torch.cuda.graph(gr, stream=s):
                stack.forward(x)
        torch.cuda.current_stream().wait_stream(s)
        ms = time_steps(gr.replay, max(args.steps // 2, 5), 3, False)
        fl, by = flops_bytes(tok, layer_shapes(m), 128)
        fl, by = fl * m["layers"], by * m["layers"]
        rec = {"tokens_per_step": tok, "ms_per_step": ms, "value": tok / (ms * 1e-3), "unit": "tokens/s",
               "path": _ext.linear_path(tok, *layer_shapes(m)[0][1:], 128)}
        if tok > 16:
            rec.update(tflops=fl / (ms * 1e-3) / 1e12, frac_of_tensor_peak=fl / (ms * 1e-3) / 1e12 / (2.0 * peaks["bf16"]),
                       bound="tmem read (tcgen05.ld of every K block's accumulator)")
        else:
            rec.update(achieved_gbs=by / (ms * 1e-3) / 1e9, frac_of_hbm_peak=by / (ms * 1e-3) / 1e9 / peaks["hbm"])
        out[name] = rec
    del stack
    torch.cuda.empty_cache()
    return out


# ------------------------------------------------------------------ CPU arm

def cpu_model_name():
    try:
        for line in open("/proc/cpuinfo"):
            if line.startswith("model name"):
                return line.split(":", 1)[1].strip()
    except OSError:
        pass
    return "unknown"


def cpu_reference(m, T, dtype, quant, group, warmup, reps):
    """The reference's CPU quantized-linear path (per_group_quant + per_group_matmul with
    group_size = K for the per-row layout; quantization.py:194-234, 265-307) as restated in
    oracle.ref_per_group_*: fp32 dequant + SGEMM on all host threads, ALL layers of the model,
    T tokens per pass.  Returns (tokens/s, measured seconds per pass [list], threads)."""
    from oracle import oracle as orc
    threads = os.cpu_count() or 1
    torch.set_num_threads(threads)          # torchrun exports OMP_NUM_THREADS=1: pin it explicitly
    g = torch.Generator().manual_seed(0)
    shapes = layer_shapes(m)
    layers = []
    for _ in range(m["layers"]):
        ws = []
        for _, N, K in shapes:
            w = (torch.randn(N, K, generator=g) * 0.02).to(torch.bfloat16)
            gs = group if quant == "per_group" else K
            ws.append((orc.ref_per_group_quant(w, gs, dtype), gs))
        layers.append(ws)
    x = torch.randn(T, m["hidden"], generator=torch.Generator().manual_seed(1)).to(torch.bfloat16)
    k_o, k_down = m["heads"] * m["hd"], m["inter"]

    def one_pass():
        h_in = x
        for (qkv, o, gu, dn) in layers:
            h = orc.ref_per_group_matmul(h_in, qkv[0][0], qkv[0][1], None, qkv[1], dtype)
            h = orc.ref_per_group_matmul(h[:, :k_o].contiguous(), o[0][0], o[0][1], None, o[1], dtype)
            h = orc.ref_per_group_matmul(h, gu[0][0], gu[0][1], None, gu[1], dtype)
            h_in = orc.ref_per_group_matmul(h[:, :k_down].contiguous(), dn[0][0], dn[0][1], None, dn[1], dtype)

    for _ in range(warmup):
        one_pass()
    ts = []
    for _ in range(reps):
        t0 = time.perf_counter()
        one_pass()
        ts.append(time.perf_counter() - t0)
    return T * len(ts) / sum(ts), ts, torch.get_num_threads()


# ------------------------------------------------------------------ main

def kernel_table(step, path, rank, dist_on, line, reps=3):
    """Per-kernel breakdown of one step (all ranks run it; rank 0 writes): kernel name, launches
    per step, mean and total microseconds, share of the summed kernel time.  Collected with the
    CUPTI activity tracer (torch.profiler) AFTER the timed runs; its times carry tracer overhead
    and are for attribution only."""
    import collections
    import torch.distributed as dist
    from torch.profiler import ProfilerActivity, profile
    if dist_on:
        dist.barrier()
    torch.cuda.synchronize()
    with profile(activities=[ProfilerActivity.CUDA]) as prof:
        for _ in range(reps):
            step()
        torch.cuda.synchronize()
    if dist_on:
        dist.barrier()
    if rank != 0:
        return
    # Kernels overlap under programmatic dependent launch (a kernel's recorded duration includes
    # its wait on the predecessor), so durations do not add up.  "exposed" partitions the step
    # instead: end_i - max(start_i, latest end so far), in start order.
    evs = [ev for ev in prof.events()
           if ev.device_type is not None and "cuda" in str(ev.device_type).lower()]
    evs.sort(key=lambda ev: ev.time_range.start)
    agg = collections.OrderedDict()
    t0 = evs[0].time_range.start if evs else 0.0
    t1 = max(ev.time_range.end for ev in evs) if evs else 0.0
    front, idle = t0, 0.0
    for ev in evs:
        tr = ev.time_range
        d = agg.setdefault(ev.name, [0, 0.0, 0.0])
        d[0] += 1
        d[1] += tr.end - tr.start
        if tr.start > front:
            idle += tr.start - front
        d[2] += max(0.0, tr.end - max(tr.start, front))
        front = max(front, tr.end)
    span = (t1 - t0) or 1.0
    with open(path, "w") as f:
        f.write(f"# {line['config']['workload']}  {line['config']['parallelism']}  ({line['config']['launch']})\n")
        f.write(f"# timed (no tracer): {line['ms_per_step']:.4f} ms/step, {line['value']:.0f} tokens/s;  under the tracer: "
                f"{span / reps / 1e3:.4f} ms/step span, {idle / reps / 1e3:.4f} ms/step idle between kernels "
                f"(rank 0, {reps} steps)\n")
        f.write("# duration = start..end incl. the wait on the predecessor under PDL (overlaps, does not add up);\n")
        f.write("# exposed = end - max(start, latest earlier end): partitions the step span\n")
        f.write("# launches/step | mean duration us | exposed us/launch | exposed us/step | share of span | kernel\n")
        for name, (n, us, ex) in sorted(agg.items(), key=lambda kv: -kv[1][2]):
            f.write(f"{n / reps:8.1f} | {us / n:9.2f} | {ex / n:9.2f} | {ex / reps:10.1f} | {ex / span:6.3f} | {name[:150]}\n")


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--model", default="0.6b", choices=list(MODELS))
    ap.add_argument("--tokens", type=int, default=4096, help="tokens per rank per step")
    ap.add_argument("--dtype", default="int8", choices=["int8", "fp8"])
    ap.add_argument("--quant", default="per_row", choices=["per_row", "per_group"])
    ap.add_argument("--group-size", type=int, default=128)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-extra-legs", action="store_true", help="skip the FP8 g=128 and model-shaped legs")
    ap.add_argument("--no-graph", action="store_true", help="launch eagerly instead of replaying a CUDA graph")
    ap.add_argument("--kernel-table", default=None, metavar="PATH",
                    help="AFTER the line is printed: replay 3 steps under the CUPTI kernel tracer and write "
                         "per-kernel-name mean durations (rank 0) -- a breakdown for profiles/, never a bench value")
    args = ap.parse_args()

    m = MODELS[args.model]
    dtype = torch.int8 if args.dtype == "int8" else torch.float8_e4m3fn
    group = args.group_size if args.quant == "per_group" else 0
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    T_rank = args.tokens
    workload = (f"qwen3-{args.model} linear stack ({4 * m['layers']} quantized linears: qkv,o,gate_up,down x "
                f"{m['layers']} layers), {args.dtype} {args.quant}"
                + (f" g={group}" if group else "") + f", {T_rank} tokens/rank/step")

    use_graph = not args.no_graph
    config = {"workload": workload, "tokens_per_step": T_rank * world, "parallelism": f"tp{world}",
              "l2": "one step streams every layer's distinct weights + activations (> 126 MB L2 "
                    "working set for the default); no explicit flush",
              "quant": args.quant, "group_size": group,
              "launch": "cuda graph replay" if use_graph else "eager"}

    # ---------------------------------------------------------------- reference (CPU) arm
    if args.impl == "reference":
        if rank != 0:
            return
        # Every step is one MEASURED pass over all layers of the model on a bounded token sample
        # (the CPU path's cost is linear in tokens); nothing is extrapolated across layers.
        T_s = min(T_rank * world, 4096)
        tok_s, ts, cores = cpu_reference(m, T_s, dtype, args.quant, group, min(max(args.warmup, 1), 2),
                                         max(args.steps, 1))
        ms_step = 1e3 * sum(ts) / len(ts)
        sample = (f"every step = all {m['layers']} layers x {T_s} of the {T_rank * world} tokens of a step, "
                  f"{len(ts)} timed steps (mean) after {min(max(args.warmup, 1), 2)} warm-up passes, torch fp32 dequant + SGEMM on {cores} threads of "
                  f"{os.cpu_count()} host CPUs ({cpu_model_name()}); activations NOT quantised "
                  "(the reference's per_group_matmul semantics)")
        line = {
            "impl": "reference", "metric": METRIC, "value": tok_s, "unit": "tokens/s", "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms_step,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32",
            "data": "synthetic (random-init N(0,0.02) bf16 weights quantised on the host, randn activations)",
            "config": config,
            "reference_arm": "reference CPU path: torch dequant + fp32 matmul (oracle.ref_per_group_matmul)",
            "cpu_baseline": {"value": tok_s, "unit": "tokens/s", "cores": cores, "kind": "port",
                             "sample": sample},
            "e2e": {"value": tok_s, "unit": "tokens/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0,
        }
        print(json.dumps(line))
        return

    # ---------------------------------------------------------------- B200 arm
    import torch.distributed as dist
    assert torch.cuda.is_available(), "bench.py --impl b200 needs a CUDA device (no CPU fallback)"
    torch.cuda.set_device(local)
    dist_on = world > 1
    if dist_on:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    from nanovllm import _ext
    assert _ext.load().ql_device_check() == 0
    peaks = load_peaks()
    workload_key = f"{args.model}/{args.dtype}/{args.quant}/{group}/T{T_rank}/tp{world}"

    stack = LinearStack(m, dtype, args.quant, group, "cuda", seed=0)
    T = T_rank * world                                     # every TP rank sees every token
    gx = torch.Generator(device="cuda").manual_seed(1)
    x_dev = torch.randn(T, m["hidden"], generator=gx, device="cuda", dtype=torch.float32).to(torch.bfloat16)
    x_host = x_dev.cpu().pin_memory()
    y_host = torch.empty((T, m["hidden"]), dtype=torch.bfloat16).pin_memory()
    x_in = torch.empty_like(x_dev)

    # The whole pass (fixed token count) is captured once into a CUDA graph and replayed, the way
    # the reference engine replays its decode step (ModelRunner.capture_cudagraph,
    # model_runner.py:316-356); --no-graph launches the 2 x 112 kernels eagerly from Python, as the
    # reference does for prefill (model_runner.py:283-284) -- that mode is CPU-launch-bound.
    graph_out = {}

    def capture(x_buf):
        s = torch.cuda.Stream()
        s.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(s):
            stack.forward(x_buf)                           # warm-up outside capture
            gr = torch.cuda.CUDAGraph()
            with torch.cuda.graph(gr, stream=s):
                out = stack.forward(x_buf)
        torch.cuda.current_stream().wait_stream(s)
        return gr, out

    if use_graph:
        graph_out["dev"] = capture(x_dev)
        graph_out["e2e"] = capture(x_in)

    def step_dev():
        if use_graph:
            graph_out["dev"][0].replay()
            return graph_out["dev"][1]
        return stack.forward(x_dev)

    # e2e: every step copies ITS input from pinned host memory and ITS result back to pinned host
    # memory.  As a serving loop would, the copies run on their own streams and are double-buffered:
    # the H2D of step i+1 and the D2H of step i-1 overlap the kernels of step i; the host reads
    # (synchronises on) the result of the previous step while the current one runs.
    cp_in, cp_out = torch.cuda.Stream(), torch.cuda.Stream()
    stage_in = [torch.empty_like(x_dev) for _ in range(2)]
    stage_out = [torch.empty_like(x_dev) for _ in range(2)]
    y_hosts = [y_host, torch.empty_like(y_host).pin_memory()]
    ev_in = [torch.cuda.Event() for _ in range(2)]
    ev_free = [torch.cuda.Event() for _ in range(2)]
    ev_done = [torch.cuda.Event() for _ in range(2)]
    ev_out = [torch.cuda.Event() for _ in range(2)]
    for e in ev_free + ev_out:
        e.record()
    e2e_i = [0]

    def step_e2e():
        i = e2e_i[0]
        b = i & 1
        e2e_i[0] = i + 1
        main = torch.cuda.current_stream()
        cp_in.wait_event(ev_free[b])                      # stage_in[b] was consumed two steps ago
        with torch.cuda.stream(cp_in):
            stage_in[b].copy_(x_host, non_blocking=True)  # H2D of this step's input
            ev_in[b].record(cp_in)
        main.wait_event(ev_in[b])
        x_in.copy_(stage_in[b])
        ev_free[b].record(main)
        if use_graph:
            graph_out["e2e"][0].replay()
            y = graph_out["e2e"][1]
        else:
            y = stack.forward(x_in)
        main.wait_event(ev_out[b])                        # stage_out[b] was read back two steps ago
        stage_out[b].copy_(y)
        ev_done[b].record(main)
        cp_out.wait_event(ev_done[b])
        with torch.cuda.stream(cp_out):
            y_hosts[b].copy_(stage_out[b], non_blocking=True)   # D2H of this step's result
            ev_out[b].record(cp_out)
        if i > 0:
            ev_out[1 - b].synchronize()                   # the caller reads the previous step's result

    def finish_e2e():
        for e in ev_out:
            torch.cuda.current_stream().wait_event(e)

    sampler = ClockSampler(local) if rank == 0 else None
    n0 = _ext.launch_count()
    stack.forward(x_dev)                                   # one eager pass counts the launches/step
    launches_per_step = _ext.launch_count() - n0
    ms = time_steps(step_dev, args.steps, args.warmup, dist_on)
    clocks = sampler.stop() if sampler else {}
    ms_e2e = time_steps(step_e2e, args.steps, max(args.warmup, 3), dist_on, finish_e2e)
    torch.cuda.synchronize()
    # the pipelined leg must have produced the real result: same input as the resident leg
    y_check = step_dev().float().cpu()
    last = (e2e_i[0] - 1) & 1
    assert torch.equal(y_hosts[last].float(), y_check), "e2e result differs from the resident-input result"

    # ---- multi-GPU correctness, visible in the line: the collective the timed path used, on one
    # row-parallel layer, against its definition (fp32 rank-order sum of the gathered bf16 partials,
    # rounded once; per_row_quant of that) -- bit for bit, prefill- and decode-sized
    tp_info = None
    if dist_on:
        tp_info = tp_check(stack, T, dtype, rank, world)

    # ---- decode-shaped leg (BASELINE.json configs[2] regime): the same stack at 16 tokens per rank
    decode = None
    if T_rank > 16:
        decode = decode_leg(stack, m, 16 * world, dtype, group, world, peaks, args, dist_on)

    # ---- BASELINE configs[3]: Qwen3-8B under tensor parallelism, in every N > 1 line of the default workload
    tp_8b = None
    if dist_on and group == 0 and dtype == torch.int8 and T_rank > 16 and args.model == "0.6b" \
            and not args.no_extra_legs:
        tp_8b = tp_8b_leg(dtype, args.quant, T_rank, rank, world, peaks, args)

    # ---- FP8 g=128 leg (the reference's per-group configuration), default workload only
    fp8_g128 = model_shaped = None
    if world == 1 and group == 0 and T_rank > 16 and args.model == "0.6b" and not args.no_extra_legs:
        model_shaped = {"prefill": model_shaped_leg(stack, m, T, dtype, args),
                        "decode": model_shaped_leg(stack, m, 16, dtype, args)}
        fp8_g128 = fp8_group_leg(m, T, peaks, args)

    # ---- dominant kernel alone (roofline): the GEMM (or GEMV for T <= 16) over all local weights
    shapes = layer_shapes(m, world)
    fl_layer, by_layer = flops_bytes(T, shapes, group)
    n_lin = 4 * m["layers"]
    acts = {}
    for (_, N, K) in shapes:
        if K not in acts:
            xa = torch.randn(T, K, device="cuda", dtype=torch.float32, generator=gx).to(torch.bfloat16)
            acts[K] = (xa,) + tuple(_ext.quant_rows(xa, dtype))

    def kernels_only():
        # the GEMM / decode kernel of every linear on pre-quantised activations (the kernel
        # ql_linear launches after its quantiser, for every T)
        for mods in stack.layers:
            for mod in mods:
                K = mod.qw.shape[1]
                _ext.gemm_w8a8(acts[K][1], acts[K][2], mod.qw, mod._sw_kernel, None, dtype, mod.group_size)

    if use_graph:
        ks = torch.cuda.Stream()
        ks.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(ks):
            kernels_only()
            kgr = torch.cuda.CUDAGraph()
            with torch.cuda.graph(kgr, stream=ks):
                kernels_only()
        torch.cuda.current_stream().wait_stream(ks)
        ms_k = time_steps(kgr.replay, max(args.steps // 2, 2), 3, False)
    else:
        ms_k = time_steps(kernels_only, max(args.steps // 2, 2), 3, False)
    k_fl = sum(2 * T * N * K for _, N, K in shapes) * m["layers"]
    k_by = sum(T * K + N * K + 4 * T + 4 * N * (K // group if group else 1) + 2 * T * N
               for _, N, K in shapes) * m["layers"]
    t_k = ms_k * 1e-3
    tensor_peak = 2.0 * peaks["bf16"]                     # 8-bit dense rate = 2 x bf16 on sm_100
    t_fl, t_by = k_fl / (tensor_peak * 1e12), k_by / (peaks["hbm"] * 1e9)
    if t_fl >= t_by:
        roof = {"bound": "tensor", "achieved": k_fl / t_k / 1e12, "peak": tensor_peak, "unit": "TFLOP/s"}
    else:
        roof = {"bound": "hbm", "achieved": k_by / t_k / 1e9, "peak": peaks["hbm"], "unit": "GB/s"}
    roof["frac"] = roof["achieved"] / roof["peak"]
    # beside the measured-peak fraction, the one against the NOMINAL datasheet peak (4.5 PFLOP/s dense
    # 8-bit, 8 TB/s HBM3e): MEASURED_PEAKS.json's bf16 burst was taken under a power-limited clock, so
    # the measured-peak fraction flatters short, full-clock launches ("smaller" item 13)
    roof["nominal_peak"] = 4500.0 if roof["bound"] == "tensor" else 8000.0
    roof["frac_nominal"] = roof["achieved"] / roof["nominal_peak"]
    roof.update(traffic=load_traffic(workload_key),
                kernel={"gemm_tcgen05_2cta": "gemm2_kernel (tcgen05 cta_group::2)",
                        "gemm_tcgen05": "gemm_kernel (tcgen05)"}.get(
                            _ext.linear_path(T, shapes[0][1], shapes[0][2], group).replace("quant+", ""),
                            "skinny_kernel (tcgen05 cluster split-K)"),
                launches=n_lin, avg_launch_us=ms_k * 1e3 / n_lin,
                tflops=k_fl / t_k / 1e12, gbs=k_by / t_k / 1e9,
                share_of_step=ms_k / ms,
                peak_source=peaks["src"] + ("; tensor peak = 2 x measured bf16 burst (8-bit dense is 2x bf16)"
                                            if roof["bound"] == "tensor" else ""))

    line = {
        "metric": METRIC, "value": T / (ms * 1e-3), "unit": "tokens/s", "n_gpus": world,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "int8" if dtype == torch.int8 else "fp8_e4m3",
        "data": "synthetic (random-init N(0,0.02) bf16 weights quantised on device, randn activations)",
        "config": config,
        "tflops": fl_layer * m["layers"] / (ms * 1e-3) / 1e12,
        "e2e": {"value": T / (ms_e2e * 1e-3), "unit": "tokens/s", "ms_per_step": ms_e2e,
                "pipeline": "double-buffered: H2D of step i+1 / D2H of step i-1 overlap the kernels of step i",
                "h2d_bytes_per_step": x_host.numel() * 2, "d2h_bytes_per_step": y_host.numel() * 2},
        "gpu_launches": int(launches_per_step) * args.steps,
        "gpu_launches_per_step": int(launches_per_step),
        "roofline": roof, "clocks": clocks,
    }
    if decode is not None:
        line["decode"] = decode
    if fp8_g128 is not None:
        line["fp8_g128"] = fp8_g128
    if model_shaped is not None:
        line["model_shaped"] = model_shaped
    if tp_8b is not None:
        line["tp_8b"] = tp_8b
    if tp_info is not None:
        line["tp_check"] = tp_info["status"]
        line["tp_collective"] = tp_info
        nmb = stack.micro_batches(T)
        config["tp_micro_batches"] = nmb
        config["tp_pipeline"] = ("2 token micro-batches on 2 streams, one NVLink collective channel each"
                                 if nmb > 1 else "single stream")
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        T_s = min(T_rank, 4096)
        tok_s, ts, cores = cpu_reference(m, T_s, dtype, args.quant, group, 1, 2)
        line["cpu_baseline"] = {"value": tok_s, "unit": "tokens/s", "cores": cores, "kind": "port",
                                "sample": f"all {m['layers']} layers x {T_s} tokens, mean of {len(ts)} measured passes "
                                          f"after 1 warm-up ({cpu_model_name()}); activations not quantised "
                                          "(reference per_group_matmul semantics)"}
    if rank == 0:
        print(json.dumps(line), flush=True)
    if args.kernel_table:
        try:
            kernel_table(step_dev, args.kernel_table, rank, dist_on, line)
        except Exception as e:  # noqa: BLE001 -- attribution only; the line above is already out
            print(f"kernel table failed: {e!r}", file=sys.stderr, flush=True)
    if dist_on:
        # The captured graphs hold NCCL kernels; tearing the communicator down under them hung
        # the first TP=2 run after the line was printed (profiles/r1o_*).  Everything is measured
        # and flushed: meet once more, then leave without the NCCL destructor.
        torch.cuda.synchronize()
        dist.barrier()
        sys.stdout.flush()
        sys.stderr.flush()
        os._exit(0)


if __name__ == "__main__":
    main()
