python -m pytest tests -m gpu -x -q 2>&1 | tail -3
python bench.py > gpurun_out/bench_v5.json 2>gpurun_out/bench_v5.err; tail -c 2500 gpurun_out/bench_v5.json
timeout 300 python tests/fuzz_gpu.py 60 2>&1 | tail -2
timeout 300 ncu --set full --clock-control none --import-source on -k regex:score_kernel -c 1 -f -o gpurun_out/score_v5 python profiles/score_only.py c3 1 > gpurun_out/ncu_v5.log 2>&1; tail -2 gpurun_out/ncu_v5.log
