#!/bin/bash
# usage: r2_quick.sh [fuzz seconds] -- the scoring / placement GPU tests, a short fuzz, one bench line
timeout 600 python -m pytest tests/test_gpu_parity.py tests/test_gpu_digests.py -x -q -m gpu -k "not reconcile" 2>&1 | tail -5
timeout 400 python tests/fuzz_gpu.py ${1:-60} 777 2>&1 | tail -4
python bench.py --steps 20 --warmup 5 > gpurun_out/r2_quick_bench.json 2> gpurun_out/r2_quick_bench.err
python - <<PY
import json
t=open('gpurun_out/r2_quick_bench.json').read(); d=json.loads(t[t.index('{"metric'):])
print('value %.4g ms %.3f e2e %.4g e2e_ms %.3f' % (d['value'], d['ms_per_step'], d['e2e']['value'], d['e2e']['ms_per_step']), 'score evals/s %.4g share %.3f' % (d['roofline']['achieved'], d['roofline']['share_of_step']), d['config']['result_sha256'][:12], 'c4 base %.4g' % d['scaling_base']['value'])
PY
