#!/bin/bash
# usage: tools/bench_scale.sh "<N list>" "<variant>;<variant>..."   (run under gpurun --gpus max(N))
IFS=';' read -ra VARS <<< "${2:---schedule overlapped --seams exact}"
for N in $1; do
 for v in "${VARS[@]}"; do
  if [ "$N" = "1" ]; then L="python bench.py"; else L="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29533 bench.py"; fi
  $L --gpus $N --steps 30 --warmup 5 --no-cpu --no-e2e $v 2>>gpurun_out/bench_scale.err | grep '^{' | tee -a gpurun_out/bench_scale.jsonl | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print('$v', 'N=', d['n_gpus'], 'ms/step', round(d['ms_per_step'],3), 'value', round(d['value']/1e9,3), 'G/s', 'kernels', {q: round(v['ms'],3) for q,v in d['roofline']['kernels'].items()}, d['clocks']['sm_mhz'], d['clocks']['reasons'])"
 done
done
