#!/bin/bash
# reconcile at scale: plain timings, then the ncu launch list of the same command
python -m pytest tests/test_gpu_parity.py -x -q -k reconcile 2>&1 | tail -3
python tests/fuzz_gpu.py 40 91000 2>&1 | tail -3
python profiles/scripts/rec_large_run.py 1e7 > gpurun_out/r2_rec_large_1e7.log 2>&1; cat gpurun_out/r2_rec_large_1e7.log
python profiles/scripts/rec_large_run.py 1e8 > gpurun_out/r2_rec_large_1e8.log 2>&1; cat gpurun_out/r2_rec_large_1e8.log
ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/r2_rec_launches_1e8.csv python profiles/scripts/rec_large_run.py 1e8 1 > gpurun_out/r2_rec_ncu_1e8.log 2>&1
python profiles/scripts/ncu_launch_summary.py gpurun_out/r2_rec_launches_1e8.csv > gpurun_out/r2_rec_launches_1e8_summary.txt 2>&1; cat gpurun_out/r2_rec_launches_1e8_summary.txt
grep -c . gpurun_out/r2_rec_launches_1e8.csv
