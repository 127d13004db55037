#!/bin/bash
# ncu --set full capture of the fused step kernel of the EfficientZero and Gumbel trees (after the plain command exited 0)
mkdir -p gpurun_out
for v in ez gmz; do
CMD="python bench.py --steps 1 --warmup 1 --no-graph --no-cpu-baseline --no-e2e --variant $v"
$CMD > gpurun_out/plain_$v.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:k_step -s 130 -c 2 -o gpurun_out/prof_step_$v -f $CMD > gpurun_out/ncu_full_$v.log 2>&1; echo ncu_$v=$?
ncu -i gpurun_out/prof_step_$v.ncu-rep --page raw --csv > gpurun_out/step_${v}_raw.csv 2>/dev/null
ncu -i gpurun_out/prof_step_$v.ncu-rep --page details --csv > gpurun_out/step_${v}_details.csv 2>/dev/null
rm -f gpurun_out/prof_step_$v.ncu-rep
done
