#!/bin/bash
# round-2 GPU call F (1 GPU): decode chain parity + default bench + 8B decode leg (+ full gpu suite with "full")
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_decode_chain.py -x -q > gpurun_out/r2f_chain_tests.log 2>&1; echo "chain tests rc=$?"
tail -15 gpurun_out/r2f_chain_tests.log
timeout 300 python bench.py --steps 20 --warmup 5 > gpurun_out/r2f_bench.json 2> gpurun_out/r2f_bench.err; echo "bench rc=$?"
tail -3 gpurun_out/r2f_bench.err
python - <<'PY'
import json
try:
    d = json.loads(open("gpurun_out/r2f_bench.json").read().strip().splitlines()[-1])
    print("value", round(d["value"]), "ms", d["ms_per_step"], "e2e", round(d["e2e"]["value"]), "roof", d["roofline"]["frac"])
    print("decode", json.dumps(d.get("decode"), indent=1))
except Exception as e:
    print("no line", e)
PY
timeout 600 python bench.py --model 8b --tokens 128 --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/r2f_bench_8b.json 2> gpurun_out/r2f_bench_8b.err; echo "bench 8b rc=$?"
tail -3 gpurun_out/r2f_bench_8b.err
python - <<'PY'
import json
try:
    d = json.loads(open("gpurun_out/r2f_bench_8b.json").read().strip().splitlines()[-1])
    print("8b decode", json.dumps(d.get("decode"), indent=1))
except Exception as e:
    print("no line", e)
PY
if [ "$1" = "full" ]; then
  timeout 1500 python -m pytest tests -m gpu -q > gpurun_out/r2f_pytest_gpu.log 2>&1; echo "gpu suite rc=$?"
  tail -8 gpurun_out/r2f_pytest_gpu.log
fi
exit 0
