#!/bin/bash
# N-GPU check (gpurun --gpus N -- 'bash tools/gpu_multi.sh N tag'): the multi-GPU tests (NCCL harness, C slab driver over
# cudaIpc, HF2D_GPUS shim) and bench.py at N through the C driver (and the NCCL harness for comparison).
N=${1:-2}; tag=${2:-m$N}
mkdir -p gpurun_out
timeout -s KILL 600 python -m pytest tests/test_zz_slab_driver_gpu.py tests/test_shim_dropin.py tests/test_gpu_parity.py -m gpu -q -p no:cacheprovider -k "slab or nccl or several_gpus" > gpurun_out/${tag}_tests.log 2>&1
echo "pytest rc=$?" >> gpurun_out/${tag}_tests.log; tail -4 gpurun_out/${tag}_tests.log
for drv in c python; do
  timeout -s KILL 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29533 bench.py --gpus $N --steps 30 --warmup 5 --no-cpu --driver $drv > gpurun_out/${tag}_bench_$drv.log 2> gpurun_out/${tag}_bench_$drv.err
  python - <<PY
import json
for l in open('gpurun_out/${tag}_bench_$drv.log'):
    if l.startswith('{'):
        d=json.loads(l); print('$drv N=', d['n_gpus'], 'ms/step', round(d['ms_per_step'],4), 'G/s', round(d['value']/1e9,3), 'e2e G/s', round(d.get('e2e',{}).get('value',0)/1e9,3), d.get('parity'), {k:round(v['ms'],3) for k,v in d['roofline']['kernels'].items()})
PY
  tail -3 gpurun_out/${tag}_bench_$drv.err
done
