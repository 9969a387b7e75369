#!/bin/bash
# ncu --set full of the plain-tile kernels of one step of the bench workload (after a plain run of the same command),
# plus the launch list of a short run.   gpurun --timeout 900 -- 'bash tools/prof_plain.sh <tag> [model]'
# The report itself can exceed what gpurun carries back: the raw and source pages are exported to (gzipped) CSV on the
# box and the report is kept only when it is small.
tag=${1:-r2}; model=${2:-keps}
mkdir -p gpurun_out
cmd="python bench.py --steps 4 --warmup 3 --no-cpu --no-e2e --model $model"
$cmd > gpurun_out/${tag}_plain.log 2>&1 || { tail -5 gpurun_out/${tag}_plain.log; exit 1; }
rep=/tmp/${tag}_prof.ncu-rep
ncu --set full --clock-control none --import-source on -k 'regex:k_(stage1_plain|stage1_tma|point|flux|step_plain)' -s 24 -c 8 -f -o ${rep%.ncu-rep} $cmd > gpurun_out/${tag}_ncu.log 2>&1
ncu -i $rep --page raw --csv > gpurun_out/${tag}_raw.csv 2>/dev/null
ncu -i $rep --page source --csv 2>/dev/null | gzip -9 > gpurun_out/${tag}_source.csv.gz
ls -la $rep gpurun_out/${tag}_raw.csv gpurun_out/${tag}_source.csv.gz
sz=$(stat -c %s $rep); if [ "$sz" -lt 30000000 ]; then cp $rep gpurun_out/; fi
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/${tag}_launches.csv $cmd > gpurun_out/${tag}_ncu_launches.log 2>&1
tail -2 gpurun_out/${tag}_plain.log | cut -c1-300; tail -3 gpurun_out/${tag}_ncu.log
