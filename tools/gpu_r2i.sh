#!/bin/bash
# round-2 GPU call I (N GPUs): the driver's scaling command at N, plus the per-kernel table
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
N=${1:-8}
export QL_AR_TIMEOUT_S=8
timeout 420 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29517 \
  bench.py --gpus $N --steps 20 --warmup 5 --kernel-table gpurun_out/r2i_ktable_n$N.txt > gpurun_out/r2i_bench_n$N.json 2> gpurun_out/r2i_bench_n$N.err
echo "bench rc=$?"
python - <<PY
import json
try:
    d=json.loads(open("gpurun_out/r2i_bench_n$N.json").read().strip().splitlines()[-1])
    print("N=$N", round(d["value"]), "tok/s", d["ms_per_step"], "ms e2e", round(d["e2e"]["value"]), d.get("tp_check"), d.get("decode", {}).get("ms_per_step"))
    print(json.dumps(d.get("tp_collective")))
except Exception as e:
    print("no line", e)
PY
tail -5 gpurun_out/r2i_bench_n$N.err
head -14 gpurun_out/r2i_ktable_n$N.txt
