#!/bin/bash
# round-2 GPU call J (N GPUs): reduce-scatter by push vs pull, one vs two micro-batches
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
N=${1:-8}
export QL_AR_TIMEOUT_S=8
run() { name=$1; shift
  env "$@" timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29517 \
    bench.py --gpus $N --steps 10 --warmup 3 > gpurun_out/r2j_n${N}_$name.json 2> gpurun_out/r2j_n${N}_$name.err
  echo "$name rc=$?"; python - <<PY
import json
try:
    d=json.loads(open("gpurun_out/r2j_n${N}_$name.json").read().strip().splitlines()[-1])
    print("  $name", round(d["value"]), "tok/s", round(d["ms_per_step"],3), "ms e2e", round(d["e2e"]["value"]), d.get("tp_check"), d["config"].get("tp_micro_batches"), [c["collective"][:30] for c in d["tp_collective"]["checked"]])
except Exception as e:
    print("  $name: no line", e)
PY
  tail -2 gpurun_out/r2j_n${N}_$name.err | cut -c1-300
}
for v in "$@"; do :; done
run pull_micro1 QL_TP_PUSH=0 QL_TP_MICRO=1
run pull_micro2 QL_TP_PUSH=0 QL_TP_MICRO=2
run push_micro2 QL_TP_PUSH=1 QL_TP_MICRO=2
