#!/bin/bash
# round 2 profiles (1 GPU): launch list of a short bench run, then ncu --set full of the three kernels
CMD="python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-scaling-base"
$CMD > gpurun_out/r2_prof_plain.log 2> gpurun_out/r2_prof_plain.err || { echo plain run failed; tail -5 gpurun_out/r2_prof_plain.err; exit 1; }
ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/r2_launches_c3.csv $CMD > gpurun_out/r2_prof_ncu1.log 2>&1
python profiles/scripts/ncu_launch_summary.py gpurun_out/r2_launches_c3.csv > gpurun_out/r2_launches_c3_summary.txt; cat gpurun_out/r2_launches_c3_summary.txt
ncu --set full --clock-control none --import-source on -k regex:score_kernel -s 14 -c 3 -o gpurun_out/r2_score_full -f $CMD > gpurun_out/r2_prof_ncu2.log 2>&1; tail -2 gpurun_out/r2_prof_ncu2.log
ncu --set full --clock-control none --import-source on -k regex:admit_coop_kernel -s 14 -c 2 -o gpurun_out/r2_admit_full -f $CMD > gpurun_out/r2_prof_ncu3.log 2>&1; tail -2 gpurun_out/r2_prof_ncu3.log
ncu --set full --clock-control none --import-source on -k regex:reconcile_small_kernel -s 2 -c 2 -o gpurun_out/r2_recsmall_full -f $CMD > gpurun_out/r2_prof_ncu4.log 2>&1; tail -2 gpurun_out/r2_prof_ncu4.log
ls -la gpurun_out/*.ncu-rep
