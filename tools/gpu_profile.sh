#!/bin/bash
# Run on the GPU box (under gpurun): plain bench, then the ncu launch list and one full capture of
# the dominant kernels of the same command.  Outputs land in gpurun_out/.
#   tools/gpu_profile.sh <tag> [bench args...]
set -u
tag=$1; shift
mkdir -p gpurun_out
CMD="python bench.py --steps 2 --warmup 3 --no-cpu-baseline $*"
$CMD > gpurun_out/plain_$tag.log 2> gpurun_out/plain_$tag.err || { echo "plain run failed"; tail -5 gpurun_out/plain_$tag.err; exit 1; }
tail -1 gpurun_out/plain_$tag.log
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -s ${SKIP:-672} -c 448 --csv \
    --log-file gpurun_out/launches_$tag.csv $CMD > gpurun_out/ncu_list_$tag.log 2>&1
echo "ncu list rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:'gemm|gemv' -s ${SKIP:-672} -c 8 \
    -o gpurun_out/prof_$tag -f $CMD > gpurun_out/ncu_full_$tag.log 2>&1
echo "ncu full rc=$?"
ls -la gpurun_out
