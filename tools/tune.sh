#!/bin/bash
# A/B of tuning builds (libhf2d_tune*.so, built with HF2D_SO_NAME / HF2D_NVCC_EXTRA="-DHF2D_ONLY_KEPS ..."): for each
# library given, the k-eps plain-tile parity test and a short bench.   gpurun -- 'bash tools/tune.sh tag libA.so libB.so ...'
tag=$1; shift
mkdir -p gpurun_out
for so in "$@"; do
  n=${so%.so}
  HF2D_SO_NAME=$so timeout -s KILL 200 python -m pytest tests/test_gpu_parity.py -m gpu -q -x -k "plain_tiles and keps" -p no:cacheprovider > gpurun_out/${tag}_${n}_test.log 2>&1
  echo "$so parity rc=$?"
  HF2D_SO_NAME=$so timeout -s KILL 200 python bench.py --steps 30 --warmup 5 --no-cpu --no-e2e > gpurun_out/${tag}_${n}_bench.log 2>&1
  python - <<PY
import json
for l in open('gpurun_out/${tag}_${n}_bench.log'):
    if l.startswith('{'):
        d=json.loads(l); print('$so', round(d['ms_per_step'],4), {k:round(v['ms'],4) for k,v in d['roofline']['kernels'].items()})
PY
done
