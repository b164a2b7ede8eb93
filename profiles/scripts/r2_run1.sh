#!/bin/bash
# round 2, first 1-GPU measurement: default bench, reconcile sweep, reconcile phase profile
python bench.py > gpurun_out/r2_bench1.json 2> gpurun_out/r2_bench1.err
tail -c 3500 gpurun_out/r2_bench1.json; tail -5 gpurun_out/r2_bench1.err
python profiles/reconcile_sweep.py 8 > gpurun_out/r2_rec_sweep1.jsonl 2> gpurun_out/r2_rec_sweep1.err
cut -c1-330 gpurun_out/r2_rec_sweep1.jsonl; tail -3 gpurun_out/r2_rec_sweep1.err
python profiles/scripts/rec_phase_prof.py > gpurun_out/r2_rec_prof1.log 2>&1; tail -20 gpurun_out/r2_rec_prof1.log
