#!/bin/bash
# what the driver runs at round end, on one GPU
python -m pytest tests/ -x -q -m gpu 2>&1 | tail -2
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -1
python bench.py --impl reference --gpus 1 --steps 20 --warmup 5 > gpurun_out/r02_bench_c3_reference_arm.json 2> /dev/null; cut -c1-200 gpurun_out/r02_bench_c3_reference_arm.json
python bench.py --gpus 1 --steps 20 --warmup 5 > gpurun_out/r02_bench_c3.json 2> gpurun_out/r02_bench_c3.err; tail -2 gpurun_out/r02_bench_c3.err
python -c "
import json; d=json.load(open('gpurun_out/r02_bench_c3.json')); r=json.load(open('gpurun_out/r02_bench_c3_reference_arm.json'))
print(d['value'], d['ms_per_step'], d['e2e']['value'], d['e2e']['ms_per_step'], d['reconcile']['ms_per_step'], d['reconcile']['e2e_ms_per_step'], d['scaling_base']['value'], 'e2e ratio', d['e2e']['value']/r['value'])"
