#!/bin/bash
# round-2 GPU call D (N GPUs): bench at N in push and NCCL transports with the per-kernel table
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
N=${1:-2}
export QL_AR_TIMEOUT_S=8
run() { # name, env...
  name=$1; shift
  env "$@" timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29517 \
    bench.py --gpus $N --steps 20 --warmup 5 --kernel-table gpurun_out/r2d_ktable_n${N}_$name.txt > gpurun_out/r2d_bench_n${N}_$name.json 2> gpurun_out/r2d_bench_n${N}_$name.err
  echo "bench $name rc=$?"; python - <<PY
import json
d=json.loads(open("gpurun_out/r2d_bench_n${N}_$name.json").read().strip().splitlines()[-1])
print("$name", round(d["value"]), d["ms_per_step"], "e2e", round(d["e2e"]["value"]), d.get("tp_check"))
PY
  head -14 gpurun_out/r2d_ktable_n${N}_$name.txt
}
run push QL_TP_PUSH=1
[ "$2" = "all" ] && run pull QL_TP_PUSH=0
run nccl QL_TP_BIG=0
