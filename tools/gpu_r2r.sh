#!/bin/bash
# stability: the driver's default scaling command at N, repeated
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
N=${1:-8}; REPS=${2:-3}
export QL_AR_TIMEOUT_S=8
for i in $(seq 1 $REPS); do
  timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29517 \
    bench.py --gpus $N --steps 20 --warmup 5 > gpurun_out/r2r_n${N}_$i.json 2> gpurun_out/r2r_n${N}_$i.err
  echo "run $i rc=$?"
  python - <<PY
import json
try:
    d=json.loads(open("gpurun_out/r2r_n${N}_$i.json").read().strip().splitlines()[-1])
    print("  N=$N", round(d["value"]), "tok/s", round(d["ms_per_step"],3), "ms e2e", round(d["e2e"]["value"]), d.get("tp_check"), "micro", d["config"].get("tp_micro_batches"), "decode", round(d["decode"]["ms_per_step"],3))
except Exception as e:
    print("  no line", e)
PY
done
