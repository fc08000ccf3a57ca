#!/bin/bash
# round-2 final single-GPU verification: full gpu suite, smoke, default bench, reference arm
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > gpurun_out/r2z_gpu.txt 2>&1
timeout 1800 python -m pytest tests -m gpu -q > gpurun_out/r2z_pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/r2z_pytest_gpu.log
tail -4 gpurun_out/r2z_pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > gpurun_out/r2z_smoke.log 2>&1; echo "smoke rc=$?"; tail -2 gpurun_out/r2z_smoke.log
timeout 600 python bench.py > gpurun_out/r2z_bench.json 2> gpurun_out/r2z_bench.err; echo "bench rc=$?"; tail -2 gpurun_out/r2z_bench.err
timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/r2z_bench_ref.json 2> gpurun_out/r2z_bench_ref.err; echo "ref rc=$?"
python - <<'PY'
import json
d = json.loads(open("gpurun_out/r2z_bench.json").read().strip().splitlines()[-1])
print("value", round(d["value"]), "ms", round(d["ms_per_step"], 4), "e2e", round(d["e2e"]["value"]), "roof", round(d["roofline"]["frac"], 4), "cpu", round(d["cpu_baseline"]["value"]))
print("decode", d["decode"]["ms_per_step"], d["decode"]["frac"], "| chain", d["decode"].get("chain", {}).get("ms_per_step"))
print("fp8_g128", json.dumps(d.get("fp8_g128")))
r = json.loads(open("gpurun_out/r2z_bench_ref.json").read().strip().splitlines()[-1])
print("ref", round(r["value"]), r["ms_per_step"], r["cpu_baseline"]["cores"])
PY
