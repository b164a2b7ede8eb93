#!/bin/bash
# usage: r2_run_ngpu.sh N [full]   -- group checks and bench on N GPUs (gpurun --gpus N)
N=$1
./tests/cpp/test_group $N > gpurun_out/r2_cpp_group_${N}gpu.log 2>&1; tail -3 gpurun_out/r2_cpp_group_${N}gpu.log
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29512 tests/run_group_ranks.py $2 > gpurun_out/r2_group_ranks_x$N.log 2>&1; grep "^\[x\|GROUP" gpurun_out/r2_group_ranks_x$N.log | tail -22
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29513 bench.py --gpus $N --steps 5 --warmup 3 > gpurun_out/r2_bench_g$N.json 2> gpurun_out/r2_bench_g$N.err
python - <<PY
import json
try:
    t=open('gpurun_out/r2_bench_g$N.json').read(); d=json.loads(t[t.index('{"metric'):])
    print('N=$N value %.4g ms %.2f e2e %.4g e2e_ms %.2f share %.3f' % (d['value'], d['ms_per_step'], d['e2e']['value'], d['e2e']['ms_per_step'], d['roofline']['share_of_step']), d['config']['result_sha256'][:12], d.get('scaling_base'))
except Exception as e:
    print('bench failed', e); print(open('gpurun_out/r2_bench_g$N.err').read()[-1500:])
PY
