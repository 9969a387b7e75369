#!/bin/bash
# quick guarded run: plain-tile parity first (a TMA byte-count mismatch would spin on its mbarrier: killed by timeout), then the suite + bench
tag=${1:-q}
mkdir -p gpurun_out
timeout -s KILL 240 python -m pytest tests/test_gpu_parity.py -m gpu -q -x -k "plain_tiles" -p no:cacheprovider > gpurun_out/${tag}_quick.log 2>&1
rc=$?; echo "quick rc=$rc" >> gpurun_out/${tag}_quick.log; tail -5 gpurun_out/${tag}_quick.log
if [ $rc -ne 0 ]; then exit 0; fi
timeout -s KILL 600 bash tools/gpu_suite.sh $tag
