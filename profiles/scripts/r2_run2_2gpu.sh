#!/bin/bash
# round 2, first multi-GPU check (gpurun --gpus 2): C++ and Python group tests, one process per GPU
# through k8sgpu_comm_init_rank, bench at N=2 with both exchanges
python -m pytest tests/test_gpu_group.py -x -q 2>&1 | tail -15 > gpurun_out/r2_group_pytest.log; cat gpurun_out/r2_group_pytest.log
./tests/cpp/test_group 2 > gpurun_out/r2_cpp_group_2gpu.log 2>&1; tail -5 gpurun_out/r2_cpp_group_2gpu.log
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 tests/run_group_ranks.py --full > gpurun_out/r2_group_ranks_x2.log 2>&1; grep -v "^W1\|^\*\*\*" gpurun_out/r2_group_ranks_x2.log | tail -25
for ex in allgather peer; do
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29513 bench.py --gpus 2 --steps 5 --warmup 3 --exchange $ex > gpurun_out/r2_bench_g2_$ex.json 2> gpurun_out/r2_bench_g2_$ex.err
tail -c 2500 gpurun_out/r2_bench_g2_$ex.json; tail -3 gpurun_out/r2_bench_g2_$ex.err
done
