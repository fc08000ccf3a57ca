#!/usr/bin/env python
"""TP decode chain: row-parallel linear -> all-reduce -> add+RMSNorm -> quant, three ways (torchrun):
  nccl    : dist.all_reduce + ql_rmsnorm_quant           (QL_PEER_AR=0 semantics)
  peer    : ql_ar_allreduce_bf16 + ql_rmsnorm_quant
  fused   : ql_ar_rmsnorm_quant (one kernel after the GEMM)
us per chain, CUDA-graph replay of 64 chains over distinct weights, max over ranks."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "nano-vllm-hw3_b200")):
    sys.path.insert(0, p)
import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402


def main():
    rank, world, local = (int(os.environ[k]) for k in ("RANK", "WORLD_SIZE", "LOCAL_RANK"))
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    from nanovllm import _comm
    from nanovllm.layers.layernorm import RMSNorm
    from nanovllm.layers.linear import RowParallelLinear
    H, I, NL = 4096, 12288, 32
    torch.set_default_dtype(torch.bfloat16)
    with torch.device("cuda"):
        layers = [RowParallelLinear(I, H) for _ in range(NL)]
        norm = RMSNorm(H)
    torch.set_default_dtype(torch.float32)
    for l in layers:
        l.weight.data.normal_(0, 0.02)
        l.quantize_(torch.int8)
    comm = _comm.get()
    for M in (1, 16, 64):
        x = torch.randn(M, I // world, device="cuda").to(torch.bfloat16)
        res = torch.randn(M, H, device="cuda").to(torch.bfloat16)

        def chain(mode):
            for l in layers:
                if mode == "fused":
                    norm.forward_quant(l.forward_partial(x), res, torch.int8, partial=True)
                elif mode == "peer":
                    norm.forward_quant(l(x), res, torch.int8)
                else:
                    y = l.forward_partial(x)
                    dist.all_reduce(y)
                    norm.forward_quant(y, res, torch.int8)
        out = {}
        for mode in ("nccl", "peer", "fused"):
            chain(mode)
            s = torch.cuda.Stream()
            s.wait_stream(torch.cuda.current_stream())
            with torch.cuda.stream(s):
                g = torch.cuda.CUDAGraph()
                with torch.cuda.graph(g, stream=s):
                    chain(mode)
            torch.cuda.current_stream().wait_stream(s)
            for _ in range(3):
                g.replay()
            dist.barrier()
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(10):
                g.replay()
            e1.record()
            torch.cuda.synchronize()
            t = torch.tensor([e0.elapsed_time(e1) * 1e3 / (10 * NL)], device="cuda")
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            out[mode] = float(t)
        if rank == 0:
            print(f"TP={world} 8B down_proj chain M={M:3d}: nccl {out['nccl']:.2f} us  peer {out['peer']:.2f} us  "
                  f"fused {out['fused']:.2f} us", flush=True)
    torch.cuda.synchronize()
    dist.barrier()
    os._exit(0)


if __name__ == "__main__":
    main()
