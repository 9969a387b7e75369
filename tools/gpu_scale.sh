#!/bin/bash
# 8-GPU box: the multi-GPU tests once, then bench.py through the C slab driver at N = 8 and N = 4 (weak scaling).
#   gpurun --gpus 8 -- 'bash tools/gpu_scale.sh tag'
tag=${1:-sc}
mkdir -p gpurun_out
timeout -s KILL 400 python -m pytest tests/test_zz_slab_driver_gpu.py tests/test_gpu_parity.py -m gpu -q -p no:cacheprovider -k "slab_driver or nccl" > gpurun_out/${tag}_tests.log 2>&1
echo "pytest rc=$?" >> gpurun_out/${tag}_tests.log; tail -3 gpurun_out/${tag}_tests.log
for N in 8 4; do
  timeout -s KILL 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29533 bench.py --gpus $N --steps 30 --warmup 5 --no-cpu > gpurun_out/${tag}_bench_n$N.log 2> gpurun_out/${tag}_bench_n$N.err
  python - <<PY
import json
for l in open('gpurun_out/${tag}_bench_n$N.log'):
    if l.startswith('{'):
        d=json.loads(l); print('N=', d['n_gpus'], 'ms/step', round(d['ms_per_step'],4), 'G/s', round(d['value']/1e9,3), 'e2e G/s', round(d.get('e2e',{}).get('value',0)/1e9,3), d.get('parity','')[:20], {k:round(v['ms'],3) for k,v in d['roofline']['kernels'].items()})
PY
  grep -i "error" gpurun_out/${tag}_bench_n$N.err | head -3
done
