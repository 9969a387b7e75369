#!/bin/bash
# usage: tools/bench2.sh N  -- N-GPU bench in the schedule/seam variants (run under gpurun --gpus N)
N=${1:-2}
for v in "--schedule overlapped --seams exact" "--schedule sequential --seams exact" "--schedule overlapped --seams lagged-chain"; do
  python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29533 bench.py --gpus $N --steps 30 --warmup 5 --no-cpu --no-e2e $v 2>&1 | tee -a gpurun_out/bench2.log | grep '^{' | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print('$v', 'N=', d['n_gpus'], 'ms/step', round(d['ms_per_step'],3), 'value', round(d['value']/1e9,3), 'G/s', 'kernels', {q: round(v['ms'],3) for q,v in d['roofline']['kernels'].items()})"
done
