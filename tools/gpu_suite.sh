#!/bin/bash
# Whole GPU suite WITHOUT -x (every test reports), then a bench line.  Usage (from /root/repo):
#   gpurun --timeout 1500 -- 'bash tools/gpu_suite.sh [tag]'
tag=${1:-r2}
mkdir -p gpurun_out
python -m pytest tests -m gpu -q -p no:cacheprovider --durations=15 > gpurun_out/${tag}_gpu_tests.log 2>&1
echo "pytest rc=$?" >> gpurun_out/${tag}_gpu_tests.log
python bench.py --steps 30 --warmup 5 > gpurun_out/${tag}_bench.log 2> gpurun_out/${tag}_bench.err
tail -40 gpurun_out/${tag}_gpu_tests.log; tail -2 gpurun_out/${tag}_bench.log
