#!/bin/bash
python -m pytest tests/test_gpu_parity.py tests/test_gang.py tests/test_gpu_digests.py -x -q -m gpu 2>&1 | tail -4
python tests/fuzz_gpu.py 60 200000 2>&1 | tail -3
K8SGPU_COOP_PROF=1 python - <<'PY' 2>&1 | grep "coop round" | head -6
import sys; sys.path.insert(0,".")
import k8sgpu_b200
from k8sgpu_b200 import capi, synth
nodes,pods=synth.gen_nodes(10000,3),synth.gen_pods(100000,3)
with capi.Context(0) as c:
    c.place_upload(nodes,pods); c.place_run()
PY
python bench.py --no-cpu-baseline > gpurun_out/r2_bench4.json 2>gpurun_out/r2_bench4.err; tail -2 gpurun_out/r2_bench4.err
python - <<'PY'
import json
d=json.load(open('gpurun_out/r2_bench4.json'))
print({k:d[k] for k in ('value','ms_per_step')}, d['e2e']['ms_per_step'], d['roofline']['avg_launch_ms'], d['roofline']['share_of_step'], d.get('scaling_base',{}).get('value'), d['config']['result_sha256'][:12])
PY
