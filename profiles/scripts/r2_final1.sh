#!/bin/bash
# round 2 final single-GPU measurements: sweeps, default bench, reference arm
python profiles/c5_sweep.py 6 > gpurun_out/r02_c5_sweep.jsonl 2> gpurun_out/r02_c5_sweep.err; tail -2 gpurun_out/r02_c5_sweep.err; grep -c '"oracle_equal": true' gpurun_out/r02_c5_sweep.jsonl; grep -c . gpurun_out/r02_c5_sweep.jsonl; grep -c 'oracle_equal": false\|oracle_equal": null' gpurun_out/r02_c5_sweep.jsonl
python profiles/reconcile_sweep.py 8 > gpurun_out/r02_reconcile_sweep.jsonl 2> gpurun_out/r02_reconcile_sweep.err; cut -c1-200 gpurun_out/r02_reconcile_sweep.jsonl
python profiles/scripts/rec_phase_prof.py > gpurun_out/r02_reconcile_c2_phases.log 2>&1; head -4 gpurun_out/r02_reconcile_c2_phases.log
python bench.py --steps 20 --warmup 5 > gpurun_out/r02_bench_c3.json 2> gpurun_out/r02_bench_c3.err; tail -2 gpurun_out/r02_bench_c3.err
python bench.py --impl reference --steps 5 --warmup 1 > gpurun_out/r02_bench_c3_reference_arm.json 2>/dev/null; cut -c1-300 gpurun_out/r02_bench_c3_reference_arm.json
python bench.py --impl reference --gpus 2 --steps 3 --warmup 1 > gpurun_out/r02_bench_c4_reference_arm.json 2>/dev/null; cut -c1-300 gpurun_out/r02_bench_c4_reference_arm.json
