#!/bin/bash
# First gpurun call of round 2: everything that was written after round 1's GPU budget ran out, newest code first,
# then the established suite and a bench line.  Usage (from /root/repo):
#   gpurun --timeout 1500 -- 'bash tools/round2_first_gpu_call.sh'
# Logs land in gpurun_out/ (merged back by gpurun).
mkdir -p gpurun_out
set -x
# 1. unrun code paths: wall distance + y+ kernels, wall-gradient write-back, time stamp, new golden groups
python -m pytest tests/test_zzz_wall_distance.py -m gpu -q                      > gpurun_out/r2_wall_distance.log 2>&1
python -m pytest tests/test_zz_reference_bodies_and_bcs_gpu.py -m gpu -q    > gpurun_out/r2_new_golden.log 2>&1
# 2. the established suite (must still be green: Stage 2 general tiles and k_pack changed)
python -m pytest tests -m gpu -q -x                                         > gpurun_out/r2_all_gpu_tests.log 2>&1
# 3. bench line (the plain-tile kernels are byte-for-byte what round 1 measured: expect ~2.8 ms/step)
python bench.py --steps 30 --warmup 5                                       > gpurun_out/r2_bench.log 2>&1
python tools/bench_wall_distance.py                                         > gpurun_out/r2_wall_distance_bench.log 2>&1
tail -3 gpurun_out/r2_wall_distance.log gpurun_out/r2_new_golden.log gpurun_out/r2_all_gpu_tests.log gpurun_out/r2_bench.log
