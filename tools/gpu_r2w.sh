#!/bin/bash
# N GPUs: everything by peer LOADS with a full-occupancy reduce kernel vs the push default
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
N=${1:-8}
export QL_AR_TIMEOUT_S=8
run() { name=$1; shift
  env "$@" timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29517 \
    bench.py --gpus $N --steps 10 --warmup 3 > gpurun_out/r2w_n${N}_$name.json 2> gpurun_out/r2w_n${N}_$name.err
  echo "$name rc=$?"; python - <<PY
import json
try:
    d=json.loads(open("gpurun_out/r2w_n${N}_$name.json").read().strip().splitlines()[-1])
    print("  $name", round(d["value"]), "tok/s", round(d["ms_per_step"],3), "ms", d.get("tp_check"))
except Exception as e:
    print("  $name: no line", e)
PY
}
run pull_all_ctas4 QL_TP_PUSH=0 QL_TP_AG_PULL=1 QL_TP_REDUCE_CTAS=4
run pull_rs_ctas4 QL_TP_PUSH=0 QL_TP_REDUCE_CTAS=4
run push_agpull_ctas4 QL_TP_AG_PULL=1 QL_TP_REDUCE_CTAS=4
