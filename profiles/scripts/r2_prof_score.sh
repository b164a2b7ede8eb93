#!/bin/bash
# ncu --set full of the scoring kernel only (round 1 and 2 of a config-3 step)
CMD="python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-scaling-base"
$CMD > gpurun_out/r2_prof_plain.log 2> gpurun_out/r2_prof_plain.err || { echo plain run failed; tail -5 gpurun_out/r2_prof_plain.err; exit 1; }
ncu --set full --clock-control none --import-source on -k regex:score_kernel -s 13 -c 2 -o gpurun_out/r2_score_full -f $CMD > gpurun_out/r2_prof_ncu2.log 2>&1; tail -2 gpurun_out/r2_prof_ncu2.log
