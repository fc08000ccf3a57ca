#!/bin/bash
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 300 python tools/chain_probe.py 0.6b 28 16 > gpurun_out/chain_probe_06b.txt 2>&1; echo rc=$?
cat gpurun_out/chain_probe_06b.txt
timeout 300 python tools/chain_probe.py 8b 8 16 > gpurun_out/chain_probe_8b.txt 2>&1; echo rc=$?
cat gpurun_out/chain_probe_8b.txt
