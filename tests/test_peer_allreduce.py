"""One-shot NVLink all-reduce (csrc/allreduce.cu) behind RowParallelLinear, on 2 GPUs.

Bar: the peer-memory all-reduce equals the fp32 rank-order sum of the gathered bf16 partials rounded
once (its definition) BIT FOR BIT at world 2, 4, 8; at world_size 2 that is also bit-identical to
NCCL's bf16 all-reduce (one addition, one rounding), at larger worlds NCCL differs by its
intermediate roundings (<= 2^-6 of max|y| checked).  Eager and CUDA-graph replay; outputs larger
than the buffer fall back to NCCL.  Needs >= 2 GPUs (skipped on the single-GPU tier; run with
`gpurun --gpus N`).
"""
import os
import socket

import pytest
import torch

pytestmark = pytest.mark.gpu


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, q):
    try:
        import sys
        root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
        for p_ in (root, os.path.join(root, "nano-vllm-hw3_b200")):
            if p_ not in sys.path:
                sys.path.insert(0, p_)
        import torch.distributed as dist
        # QL_TP_BIG=0: this file is about the one-shot (decode-sized) kernels; outputs beyond their
        # buffer must take NCCL here (the prefill-sized path has tests/test_tp_collective.py)
        os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), QL_TP_BIG="0")
        torch.cuda.set_device(rank)
        dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", rank))
        from nanovllm import _comm, _ext
        from nanovllm.layers.linear import RowParallelLinear
        torch.manual_seed(7)                                     # same full weight on both ranks
        prev = torch.get_default_dtype()
        torch.set_default_dtype(torch.bfloat16)
        with torch.device("cuda"):
            layer = RowParallelLinear(3072, 1024, bias=True)
        torch.set_default_dtype(prev)
        w_full = (torch.randn(1024, 3072) * 0.02).to(torch.bfloat16)
        layer.weight.weight_loader(layer.weight, w_full.cuda())
        layer.bias.weight_loader(layer.bias, torch.randn(1024).to(torch.bfloat16).cuda())
        layer.quantize_(torch.int8)
        comm = _comm.get()
        assert comm is not None and comm.world == world
        n0 = _ext.launch_count()
        for M in (1, 16, 48, 256):
            x = torch.randn(M, 3072 // world, device="cuda").to(torch.bfloat16)
            y = layer(x)                                          # peer-memory path
            part = layer._linear(x, layer.bias if rank == 0 else None)
            parts = [torch.empty_like(part) for _ in range(world)]
            dist.all_gather(parts, part)
            acc = parts[0].float()
            for t_ in parts[1:]:
                acc = acc + t_.float()                            # fp32, rank order: the kernel's definition
            assert torch.equal(y, acc.to(torch.bfloat16)), f"M={M}: peer all-reduce != fp32 rank-order sum"
            dist.all_reduce(part)                                 # reference path (linear.py:150-152)
            if world == 2:
                assert torch.equal(y, part), f"M={M}: peer all-reduce != NCCL"
            else:                                                 # NCCL rounds between ring steps
                assert (y.float() - part.float()).abs().max() <= 2.0 ** -6 * part.float().abs().max()
        assert _ext.launch_count() > n0
        # CUDA-graph replay (the engine captures its decode step, model_runner.py:316-356)
        x = torch.randn(16, 3072 // world, device="cuda").to(torch.bfloat16)
        ref = layer(x).clone()
        s = torch.cuda.Stream()
        s.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(s):
            g = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g, stream=s):
                y1 = layer(x)
                y2 = layer(x)                                    # two all-reduces in one graph
        torch.cuda.current_stream().wait_stream(s)
        for _ in range(5):
            g.replay()
        torch.cuda.synchronize()
        assert torch.equal(y1, ref) and torch.equal(y2, ref)
        # fused all-reduce + add + RMSNorm + quant == the three separate steps, bit for bit
        from nanovllm.layers.layernorm import RMSNorm
        norm = RMSNorm(1024).cuda().to(torch.bfloat16)
        norm.weight.data.normal_(1.0, 0.1, generator=torch.Generator(device="cuda").manual_seed(3))
        for M in (1, 16, 130):
            x = torch.randn(M, 3072 // world, device="cuda").to(torch.bfloat16)
            resid = torch.randn(M, 1024, generator=torch.Generator().manual_seed(M)).to(torch.bfloat16).cuda()
            qa_ref, res_ref = norm.forward_quant(layer(x), resid, torch.int8)
            qa, res = norm.forward_quant(layer.forward_partial(x), resid, torch.int8, partial=True)
            assert torch.equal(res, res_ref) and torch.equal(qa.q, qa_ref.q) and torch.equal(qa.s, qa_ref.s), M
        gq = torch.cuda.CUDAGraph()
        xg = torch.randn(16, 3072 // world, device="cuda").to(torch.bfloat16)
        rg = torch.randn(16, 1024, generator=torch.Generator().manual_seed(99)).to(torch.bfloat16).cuda()
        qa_ref, res_ref = norm.forward_quant(layer(xg), rg, torch.int8)
        s2 = torch.cuda.Stream()
        s2.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(s2):
            with torch.cuda.graph(gq, stream=s2):
                qa_g, res_g = norm.forward_quant(layer.forward_partial(xg), rg, torch.int8, partial=True)
                y_after = layer(xg)                               # plain all-reduce after the fused one
        torch.cuda.current_stream().wait_stream(s2)
        for _ in range(4):
            gq.replay()
        torch.cuda.synchronize()
        assert torch.equal(qa_g.q, qa_ref.q) and torch.equal(res_g, res_ref) and torch.equal(y_after, layer(xg))
        # larger than the symmetric buffer -> NCCL fallback still correct
        xb = torch.randn(2048, 3072 // world, device="cuda").to(torch.bfloat16)
        assert not comm.fits(2048 * 1024)
        yb = layer(xb)
        part = layer._linear(xb, layer.bias if rank == 0 else None)
        dist.all_reduce(part)
        assert torch.equal(yb, part)                              # both went through NCCL
        torch.cuda.synchronize()
        dist.barrier()
        q.put((rank, "ok"))
    except Exception as e:  # noqa: BLE001
        import traceback
        q.put((rank, "FAIL " + repr(e) + "\n" + traceback.format_exc()))
    finally:
        q.close()
        q.join_thread()                       # flush the result before leaving without destructors
        os._exit(0)


@pytest.mark.parametrize("world", [2, 4, 8])
def test_peer_allreduce_matches_reference_sum(world):
    if torch.cuda.device_count() < world:
        pytest.skip(f"needs {world} GPUs")
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, q)) for r in range(world)]
    for p in procs:
        p.start()
    res = [q.get(timeout=120) for _ in procs]
    for p in procs:
        p.join(timeout=30)
    for rank, msg in res:
        assert msg == "ok", f"rank {rank}: {msg}"
