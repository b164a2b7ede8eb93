#!/bin/bash
# final round-2 single-GPU batch: full GPU suite, sweeps, bench, reference arm, ncu captures
python -m pytest tests -m gpu -x -q 2>&1 | tail -3 | tee gpurun_out/r02_pytest_gpu.log
python profiles/c5_sweep.py 6 > gpurun_out/r02_c5_sweep.jsonl 2> gpurun_out/r02_c5_sweep.err; grep -c '"oracle_equal": true' gpurun_out/r02_c5_sweep.jsonl; grep -c . gpurun_out/r02_c5_sweep.jsonl
python bench.py --steps 20 --warmup 5 > gpurun_out/r02_bench_c3.json 2> gpurun_out/r02_bench_c3.err; tail -2 gpurun_out/r02_bench_c3.err
python bench.py --impl reference --steps 5 --warmup 1 > gpurun_out/r02_bench_c3_reference_arm.json 2>/dev/null
bash profiles/scripts/r2_prof.sh > gpurun_out/r02_prof.log 2>&1; tail -5 gpurun_out/r02_prof.log
python profiles/scripts/c4_prof.py 2>&1 | grep -E "^order|round 0 |round 5 " | head -6
