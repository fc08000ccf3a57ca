#!/bin/bash
# bench variants smoke: every CLI combination a user might run must produce a line
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
run() { name=$1; shift
  timeout 900 python bench.py --no-cpu-baseline --steps 5 --warmup 3 "$@" > gpurun_out/r2s_$name.json 2> gpurun_out/r2s_$name.err; rc=$?
  python - "$name" $rc <<'PY'
import json, sys
name, rc = sys.argv[1], sys.argv[2]
try:
    d = json.loads(open(f"gpurun_out/r2s_{name}.json").read().strip().splitlines()[-1])
    dec = d.get("decode") or {}
    print(f"{name:22s} rc={rc} {d['value']:12.0f} tok/s {d['ms_per_step']:8.3f} ms  roof {d['roofline']['bound']} {d['roofline']['frac']:.3f}  e2e {d['e2e']['value']:.0f}  decode {dec.get('ms_per_step')} {dec.get('frac')}")
except Exception as e:
    print(f"{name:22s} rc={rc} NO LINE {e}"); print(open(f"gpurun_out/r2s_{name}.err").read()[-600:])
PY
}
run fp8_row --dtype fp8
run i8_g128 --quant per_group
run f8_g128 --dtype fp8 --quant per_group
run i8_t16 --tokens 16
run i8_t256 --tokens 256
run eager --no-graph
run m17 --model 1.7b
run m8b_t8192 --model 8b --tokens 8192
run m8b_f8g128 --model 8b --tokens 4096 --dtype fp8 --quant per_group
