#!/bin/bash
# round-2 GPU call N (N GPUs): all-gather of codes by peer loads (QL_TP_AG_PULL=1) vs posted stores (0)
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
N=${1:-8}
export QL_AR_TIMEOUT_S=8
run() { name=$1; shift
  env "$@" timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29517 \
    bench.py --gpus $N --steps 10 --warmup 3 > gpurun_out/r2n_n${N}_$name.json 2> gpurun_out/r2n_n${N}_$name.err
  echo "$name rc=$?"; python - <<PY
import json
try:
    d=json.loads(open("gpurun_out/r2n_n${N}_$name.json").read().strip().splitlines()[-1])
    print("  $name", round(d["value"]), "tok/s", round(d["ms_per_step"],3), "ms e2e", round(d["e2e"]["value"]), d.get("tp_check"), d["config"].get("tp_micro_batches"))
except Exception as e:
    print("  $name: no line", e)
PY
  grep -v "OMP_NUM\|\*\*\*\*" gpurun_out/r2n_n${N}_$name.err | tail -2 | cut -c1-300
}
run ag_pull QL_TP_AG_PULL=1
run ag_push QL_TP_AG_PULL=0
[ -n "$2" ] && run ag_pull_ctas2 QL_TP_AG_PULL=1 QL_TP_REDUCE_CTAS=2
if [ "$N" != "8" ]; then run ag_pull_micro2 QL_TP_AG_PULL=1 QL_TP_MICRO=2; fi
exit 0
