#!/bin/bash
# round-2 GPU call G (1 GPU): decode leg with the L2 weight prefetch, several prefetch sizes
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
show() { python - "$1" <<'PY'
import json, sys
try:
    d = json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])["decode"]
    for k in ("per_linear_launches", "per_linear_launches_l2_prefetch", "chain"):
        if k in d:
            print(f"  {k:34s} {d[k]['ms_per_step']:.4f} ms  frac {d[k]['frac']:.3f}  launches {d[k]['gpu_launches_per_step']}")
    print("  headline", d["ms_per_step"], d["frac"], d["kernel"])
except Exception as e:
    print("no line", e); print(open(sys.argv[1].replace(".json", ".err")).read()[-1500:])
PY
}
for mb in ${MBS:-16 48 96}; do
  echo "== 8b prefetch ${mb} MB"
  QL_DECODE_PREFETCH_MB=$mb timeout 600 python bench.py --model 8b --tokens 128 --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/r2g_8b_pf$mb.json 2> gpurun_out/r2g_8b_pf$mb.err
  show gpurun_out/r2g_8b_pf$mb.json
done
echo "== 0.6b default"
timeout 300 python bench.py --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/r2g_06b.json 2> gpurun_out/r2g_06b.err
show gpurun_out/r2g_06b.json
