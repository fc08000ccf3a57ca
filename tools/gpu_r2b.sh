#!/bin/bash
# round-2 GPU call B (1 GPU): emulated TP collective, model parity, full suite, bench
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_tp_emulated.py -x -q > gpurun_out/r2b_tp_emulated.log 2>&1; echo "tp_emulated rc=$?" | tee -a gpurun_out/r2b_tp_emulated.log
tail -15 gpurun_out/r2b_tp_emulated.log
timeout 900 python -m pytest tests/test_model_parity.py -x -q > gpurun_out/r2b_model_parity.log 2>&1; echo "model_parity rc=$?" | tee -a gpurun_out/r2b_model_parity.log
tail -5 gpurun_out/r2b_model_parity.log
timeout 1500 python -m pytest tests -m gpu -q > gpurun_out/r2b_pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/r2b_pytest_gpu.log
tail -8 gpurun_out/r2b_pytest_gpu.log
timeout 600 python bench.py > gpurun_out/r2b_bench.json 2> gpurun_out/r2b_bench.err; echo "bench rc=$?"
tail -c 1500 gpurun_out/r2b_bench.json
