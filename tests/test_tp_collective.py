"""Prefill-sized row-parallel outputs across REAL GPUs: GEMM-epilogue reduce-scatter + owner-side
quantiser + all-gather (csrc/tp_collective.cu, gemm2_push_kernel) behind
``RowParallelLinear.forward_quant`` / ``.forward``, against the definition

    y = bf16( fp32 rank-order sum of the gathered bf16 partials );  (q, s) = per_row_quant(y)

bit for bit, in both transports (push from the GEMM epilogue, pull by the owner), for M with and
without tile tails, eagerly and under CUDA-graph replay, and against NCCL's all-reduce
(reference linear.py:150-152; identical at world 2, within bf16 rounding above).  Needs >= 2 GPUs
(skipped on the single-GPU tier, where tests/test_tp_emulated.py runs the same kernels; the
driver-visible multi-GPU check is bench.py's ``tp_check``).
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
        os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), QL_TP_TOKENS="8192",
                          QL_TP_HIDDEN="2048")
        torch.cuda.set_device(rank)
        dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", rank))
        from nanovllm import _comm, _ext
        from nanovllm.layers.linear import MergedColumnParallelLinear, RowParallelLinear
        prev = torch.get_default_dtype()
        torch.set_default_dtype(torch.bfloat16)
        with torch.device("cuda"):
            layer = RowParallelLinear(2048, 1024, bias=True)
            grouped = RowParallelLinear(2048, 1024)
            nxt = MergedColumnParallelLinear(1024, [512 * world, 512 * world])
        torch.set_default_dtype(prev)
        g = torch.Generator().manual_seed(7)                     # same full weights on every rank
        w_full = (torch.randn(1024, 2048, generator=g) * 0.02).to(torch.bfloat16).cuda()
        layer.weight.weight_loader(layer.weight, w_full)
        layer.bias.weight_loader(layer.bias, torch.randn(1024, generator=g).to(torch.bfloat16).cuda())
        grouped.weight.weight_loader(grouped.weight, w_full)
        nxt.weight.data.copy_((torch.randn(nxt.weight.shape, generator=g) * 0.02).to(torch.bfloat16))
        layer.quantize_(torch.int8)
        grouped.quantize_(torch.float8_e4m3fn, "per_group", 128)
        nxt.quantize_(torch.int8)
        comm = _comm.get()
        assert comm is not None and comm.world == world and comm.tp_tokens == 8192

        def definition(lin, x):
            part = lin.forward_partial(x)
            parts = [torch.empty_like(part) for _ in range(world)]
            dist.all_gather(parts, part)
            acc = parts[0].float()
            for t_ in parts[1:]:
                acc = acc + t_.float()
            return part, acc.to(torch.bfloat16)

        n0 = _ext.launch_count()
        for push in (True, False):
            comm.push = push
            for M in (4096, 1000, 777, 600):
                x = torch.randn(M, 2048 // world, generator=torch.Generator().manual_seed(M)).to(torch.bfloat16).cuda()
                assert comm.fits_big(M, 1024) and not comm.fits(M * 1024)
                for lin, dt in ((layer, torch.int8), (grouped, torch.float8_e4m3fn)):
                    part, y_def = definition(lin, x)
                    qa = lin.forward_quant(x)                             # reduce-scatter + quant + all-gather
                    q_def, s_def = _ext.quant_rows(y_def, dt)
                    assert torch.equal(qa.s, s_def), (push, M, dt)
                    assert torch.equal(qa.q.view(torch.uint8), q_def.view(torch.uint8)), (push, M, dt)
                    y = lin(x)                                            # bf16 form: all-reduce semantics
                    assert torch.equal(y, y_def), (push, M, dt)
                    dist.all_reduce(part)                                 # NCCL, the reference path
                    if world == 2:
                        assert torch.equal(y, part)
                    else:
                        assert (y.float() - part.float()).abs().max() <= 2.0 ** -6 * part.float().abs().max()
                # hand-off into the next column-parallel linear: same result as the unfused chain
                qa = layer.forward_quant(x)
                h = nxt(qa)
                _, y_def = definition(layer, x)
                assert torch.equal(h, nxt(y_def)), (push, M)
        assert _ext.launch_count() > n0
        # CUDA-graph replay of two chained collectives
        comm.push = True
        x = torch.randn(1024, 2048 // world, generator=torch.Generator().manual_seed(5)).to(torch.bfloat16).cuda()
        _, y_def = definition(layer, x)
        q_def, s_def = _ext.quant_rows(y_def, torch.int8)
        st = torch.cuda.Stream()
        st.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(st):
            gr = torch.cuda.CUDAGraph()
            with torch.cuda.graph(gr, stream=st):
                y1 = layer(x)
                qa = layer.forward_quant(x)
                h = nxt(qa)
        torch.cuda.current_stream().wait_stream(st)
        h_def = nxt(y_def)
        for _ in range(4):
            gr.replay()
        torch.cuda.synchronize()
        assert torch.equal(y1, y_def) and torch.equal(qa.q, q_def) and torch.equal(qa.s, s_def)
        assert torch.equal(h, h_def)
        # larger than the symmetric buffers -> NCCL fallback still right
        xb = torch.randn(8200, 2048 // world, generator=torch.Generator().manual_seed(9)).to(torch.bfloat16).cuda()
        assert not comm.fits_big(8200, 1024)
        part, y_def = definition(layer, xb)
        yb = layer(xb)
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
        q.join_thread()
        os._exit(0)


@pytest.mark.parametrize("world", [2, 4, 8])
def test_tp_collective_matches_definition(world):
    if torch.cuda.device_count() < world:
        pytest.skip(f"needs {world} GPUs")
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, q)) for r in range(world)]
    for p in procs:
        p.start()
    res = [q.get(timeout=240) for _ in procs]
    for p in procs:
        p.join(timeout=30)
    for rank, msg in res:
        assert msg == "ok", f"rank {rank}: {msg}"
