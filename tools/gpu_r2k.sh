#!/bin/bash
# round-2 GPU call K (1 GPU): per-group CTA-pair kernel: parity tests, then the per-group sweep with
# the pair kernel and (QL_GROUPED_PAIR=0) the single-CTA kernel
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_parity.py -x -q -k "groups" > gpurun_out/r2k_tests.log 2>&1; echo "tests rc=$?"
tail -6 gpurun_out/r2k_tests.log
for mode in 1 0; do
  QL_GROUPED_PAIR=$mode timeout 600 python benchmarks/sweep.py --models ${MODELS:-0.6b,8b} --M ${MS:-1024,4096,8192} --modes f8_g128,i8_g128 \
    --out gpurun_out/r2k_sweep_pair$mode.json > gpurun_out/r2k_sweep_pair$mode.log 2>&1; echo "sweep pair=$mode rc=$?"
  cat gpurun_out/r2k_sweep_pair$mode.log | tail -50
done
