#!/bin/bash
# round-2 GPU call C (2 GPUs): the prefill-sized TP collective across real GPUs, then bench at N=2
# in its three transports (push from the GEMM epilogue / pull by the owner / NCCL all-reduce).
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
N=${1:-2}
export QL_AR_TIMEOUT_S=8
nvidia-smi topo -m > gpurun_out/r2c_topo.txt 2>&1
timeout 600 python -m pytest tests/test_tp_collective.py -x -q > gpurun_out/r2c_tp_collective.log 2>&1; echo "tp_collective rc=$?" | tee -a gpurun_out/r2c_tp_collective.log
tail -30 gpurun_out/r2c_tp_collective.log
timeout 600 python -m pytest tests/test_peer_allreduce.py -x -q > gpurun_out/r2c_peer_ar.log 2>&1; echo "peer_ar rc=$?" | tee -a gpurun_out/r2c_peer_ar.log
tail -5 gpurun_out/r2c_peer_ar.log
run() { # name, env...
  name=$1; shift
  env "$@" timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29517 \
    bench.py --gpus $N --steps 20 --warmup 5 > gpurun_out/r2c_bench_n${N}_$name.json 2> gpurun_out/r2c_bench_n${N}_$name.err
  echo "bench $name rc=$?"; tail -c 900 gpurun_out/r2c_bench_n${N}_$name.json; echo
}
run push QL_TP_PUSH=1
run pull QL_TP_PUSH=0
run nccl QL_TP_BIG=0
