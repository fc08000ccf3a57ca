#!/bin/bash
# round-2 GPU call A: full gpu test suite, smoke, bench (graph + eager), memcheck on the reduced selection
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > gpurun_out/r2a_gpu.txt 2>&1
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/r2a_pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/r2a_pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r2a_smoke.log 2>&1; echo "smoke rc=$?" | tee -a gpurun_out/r2a_smoke.log
timeout 600 python bench.py > gpurun_out/r2a_bench.json 2> gpurun_out/r2a_bench.err; echo "bench rc=$?"
timeout 300 python bench.py --no-graph --no-cpu-baseline > gpurun_out/r2a_bench_eager.json 2> gpurun_out/r2a_bench_eager.err; echo "bench eager rc=$?"
timeout 300 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/r2a_bench_ref.json 2>&1; echo "ref rc=$?"
timeout 900 compute-sanitizer --tool memcheck --error-exitcode 3 python tools/sanitize_cases.py > gpurun_out/r2a_memcheck.log 2>&1; echo "memcheck rc=$?" | tee -a gpurun_out/r2a_memcheck.log
tail -5 gpurun_out/r2a_pytest_gpu.log
tail -3 gpurun_out/r2a_memcheck.log
