#!/bin/bash
# final round-2 single-GPU batch: full GPU suite, smoke, both bench arms, config-5 sweep, ncu captures
python -m pytest tests -m gpu -x -q 2>&1 | tail -3 | tee gpurun_out/r02_pytest_gpu.log
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -1
python bench.py --impl reference --gpus 1 --steps 20 --warmup 5 > gpurun_out/r02_bench_c3_reference_arm.json 2> /dev/null; cut -c1-160 gpurun_out/r02_bench_c3_reference_arm.json
python bench.py --gpus 1 --steps 20 --warmup 5 > gpurun_out/r02_bench_c3.json 2> gpurun_out/r02_bench_c3.err; tail -2 gpurun_out/r02_bench_c3.err
python -c "
import json; d=json.load(open('gpurun_out/r02_bench_c3.json')); r=json.load(open('gpurun_out/r02_bench_c3_reference_arm.json'))
print(d['value'], d['ms_per_step'], d['e2e']['value'], d['e2e']['ms_per_step'], d['reconcile']['ms_per_step'], d['reconcile']['e2e_ms_per_step'], d['scaling_base']['value'], 'e2e ratio', d['e2e']['value']/r['value'], d['roofline'])"
python profiles/c5_sweep.py 6 > gpurun_out/r02_c5_sweep.jsonl 2> gpurun_out/r02_c5_sweep.err; grep -c '"oracle_equal": true' gpurun_out/r02_c5_sweep.jsonl; grep -c . gpurun_out/r02_c5_sweep.jsonl
bash profiles/scripts/r2_prof.sh > gpurun_out/r02_prof.log 2>&1; tail -5 gpurun_out/r02_prof.log
python profiles/scripts/admit_phases.py 2>&1 | tail -15 > gpurun_out/r02_admit_phases.log
python tests/fuzz_gpu.py 150 4242 2>&1 | tail -3
