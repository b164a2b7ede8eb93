#!/bin/bash
python -m pytest tests -m gpu -x -q 2>&1 | tail -6
python tests/fuzz_gpu.py 60 123000 2>&1 | tail -3
python bench.py > gpurun_out/r2_bench2.json 2> gpurun_out/r2_bench2.err; tail -3 gpurun_out/r2_bench2.err
python - <<'PY'
import json
d=json.load(open('gpurun_out/r2_bench2.json'))
print({k:d[k] for k in ('value','ms_per_step')}, d['e2e']['ms_per_step'], d['roofline']['achieved'], d['roofline']['avg_launch_ms'], d['roofline']['share_of_step'], d.get('scaling_base',{}).get('value'), d['config']['result_sha256'][:12], d['reconcile']['ms_per_step'], d['reconcile']['e2e_ms_per_step'])
PY
