#!/bin/bash
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_decode_chain.py -x -q 2>&1 | tail -15
QL_DEBUG=1 timeout 300 python tools/chain_probe.py 0.6b 28 16 > gpurun_out/chain_probe_06b.txt 2>&1; echo rc=$?
cat gpurun_out/chain_probe_06b.txt
timeout 300 python tools/chain_probe.py 8b 8 16 > gpurun_out/chain_probe_8b.txt 2>&1; echo rc=$?
cat gpurun_out/chain_probe_8b.txt
