#!/bin/bash
# round-2 GPU call H (1 GPU): compute-sanitizer memcheck / racecheck / synccheck on the reduced case selection;
# eager (no graph) bench line
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
for tool in ${SAN_TOOLS:-}; do
  timeout 1200 compute-sanitizer --tool $tool --error-exitcode 3 python tools/sanitize_cases.py > gpurun_out/r2h_$tool.log 2>&1
  echo "$tool rc=$?" | tee -a gpurun_out/r2h_$tool.log
  grep -E "ERROR SUMMARY|RACECHECK SUMMARY|FAIL|ALL OK|Error|error" gpurun_out/r2h_$tool.log | head -8
done
timeout 300 python bench.py --no-graph --no-cpu-baseline --steps 10 --warmup 3 > gpurun_out/r2h_bench_eager.json 2> gpurun_out/r2h_bench_eager.err; echo "bench eager rc=$?"
python - <<'PY'
import json
try:
    d = json.loads(open("gpurun_out/r2h_bench_eager.json").read().strip().splitlines()[-1])
    print("eager", round(d["value"]), "tok/s", d["ms_per_step"], "ms; e2e", round(d["e2e"]["value"]))
except Exception as e:
    print("no line", e)
PY
