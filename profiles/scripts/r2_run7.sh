#!/bin/bash
python profiles/score_only.py c4s 5 2>&1 | tail -1
python -m pytest tests/test_gpu_parity.py -x -q 2>&1 | tail -2
python profiles/scripts/c4_prof.py 2>&1 | grep "^order"
python bench.py --no-cpu-baseline --no-reconcile > gpurun_out/r2_bench3.json 2>gpurun_out/r2_bench3.err
python - <<'PY'
import json
d=json.load(open('gpurun_out/r2_bench3.json'))
print({k:d[k] for k in ('value','ms_per_step')}, d['e2e']['ms_per_step'], d['roofline']['avg_launch_ms'], d.get('scaling_base',{}).get('value'))
PY
