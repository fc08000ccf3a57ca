#!/bin/bash
# round-2 GPU call E (N GPUs): two-channel micro-batch pipeline vs single stream vs NCCL
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
N=${1:-2}
export QL_AR_TIMEOUT_S=8
if [ "$2" = "tests" ]; then
  timeout 900 python -m pytest tests/test_tp_emulated.py -x -q > gpurun_out/r2e_tp_emulated.log 2>&1; echo "tp_emulated rc=$?" | tee -a gpurun_out/r2e_tp_emulated.log
  tail -5 gpurun_out/r2e_tp_emulated.log
  timeout 600 python -m pytest tests/test_tp_collective.py tests/test_peer_allreduce.py -x -q > gpurun_out/r2e_tp_collective.log 2>&1; echo "tp_collective rc=$?" | tee -a gpurun_out/r2e_tp_collective.log
  tail -5 gpurun_out/r2e_tp_collective.log
fi
run() { # name, env...
  name=$1; shift
  env "$@" timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29517 \
    bench.py --gpus $N --steps 20 --warmup 5 --kernel-table gpurun_out/r2e_ktable_n${N}_$name.txt > gpurun_out/r2e_bench_n${N}_$name.json 2> gpurun_out/r2e_bench_n${N}_$name.err
  echo "bench $name rc=$?"; python - <<PY
import json
try:
    d=json.loads(open("gpurun_out/r2e_bench_n${N}_$name.json").read().strip().splitlines()[-1])
    print("$name", round(d["value"]), d["ms_per_step"], "e2e", round(d["e2e"]["value"]), d.get("tp_check"), d["config"].get("tp_micro_batches"))
except Exception as e:
    print("$name: no line", e)
PY
  tail -3 gpurun_out/r2e_bench_n${N}_$name.err
  head -12 gpurun_out/r2e_ktable_n${N}_$name.txt
}
run micro2 QL_TP_MICRO=2
run micro1 QL_TP_MICRO=1
[ "$3" = "nccl" ] && run nccl QL_TP_BIG=0
exit 0
