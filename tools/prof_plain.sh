#!/bin/bash
# ncu --set full of the two plain-tile kernels of one step of the bench workload (after a plain run of the same command).
#   gpurun --timeout 900 -- 'bash tools/prof_plain.sh <tag> [model]'
tag=${1:-r2}; model=${2:-keps}
mkdir -p gpurun_out
cmd="python bench.py --steps 4 --warmup 3 --no-cpu --no-e2e --model $model"
$cmd > gpurun_out/${tag}_plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k 'regex:k_(stage1_plain|point|flux)' -s 20 -c 5 -f -o gpurun_out/${tag}_prof $cmd > gpurun_out/${tag}_ncu.log 2>&1
tail -2 gpurun_out/${tag}_plain.log | cut -c1-600; tail -5 gpurun_out/${tag}_ncu.log
