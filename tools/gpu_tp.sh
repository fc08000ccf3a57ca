#!/bin/bash
# Multi-GPU check of the tensor-parallel path on N GPUs of one box (gpurun --gpus N -- tools/gpu_tp.sh N [tag]):
# the two multi-GPU pytest files (skipped by the driver's 1-GPU tier), then bench.py at N.
# Logs land in gpurun_out/<tag>_*; copy the ones worth keeping into profiles/.
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
N=${1:-2}
TAG=${2:-r3tp}
export QL_AR_TIMEOUT_S=8
for t in tp_collective peer_allreduce; do
  timeout 600 python -m pytest tests/test_$t.py -x -q > gpurun_out/${TAG}_n${N}_$t.log 2>&1
  echo "$t rc=$?" | tee -a gpurun_out/${TAG}_n${N}_$t.log
  tail -4 gpurun_out/${TAG}_n${N}_$t.log
done
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29517 \
  bench.py --gpus $N --steps 20 --warmup 5 > gpurun_out/${TAG}_bench_n${N}.json 2> gpurun_out/${TAG}_bench_n${N}.err
echo "bench rc=$?"; tail -c 1200 gpurun_out/${TAG}_bench_n${N}.json; echo; tail -5 gpurun_out/${TAG}_bench_n${N}.err
