"""One-shot NVLink all-reduce (csrc/allreduce.cu) behind RowParallelLinear, on 2 GPUs.

Bar: at world_size 2 the peer-memory all-reduce (fp32 sum of the two bf16 partials, one rounding)
is BIT-IDENTICAL to NCCL's bf16 all-reduce of the same partials, eagerly and replayed from a CUDA
graph; outputs larger than the symmetric buffer fall back to NCCL.  Needs >= 2 GPUs (skipped on
the single-GPU tier; run with `gpurun --gpus 2`).
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
        os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
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
            dist.all_reduce(part)                                 # reference path (linear.py:150-152)
            assert torch.equal(y, part), f"M={M}: peer all-reduce != NCCL"
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
        # larger than the symmetric buffer -> NCCL fallback still correct
        xb = torch.randn(2048, 3072 // world, device="cuda").to(torch.bfloat16)
        assert not comm.fits(2048 * 1024)
        yb = layer(xb)
        part = layer._linear(xb, layer.bias if rank == 0 else None)
        dist.all_reduce(part)
        assert torch.equal(yb, part)
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


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs 2 GPUs")
def test_peer_allreduce_matches_nccl_tp2():
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = [q.get(timeout=90) for _ in procs]
    for p in procs:
        p.join(timeout=30)
    for rank, msg in res:
        assert msg == "ok", f"rank {rank}: {msg}"
