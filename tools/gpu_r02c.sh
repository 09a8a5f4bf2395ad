#!/bin/bash
# round 2, multi-GPU pass (run with gpurun --gpus N): the multi-GPU tests, the C++ program on every GPU, the bench under torchrun
set -x
N=${1:-2}
mkdir -p gpurun_out
nvidia-smi -L > gpurun_out/r02c_gpus_n$N.txt
timeout 900 python -m pytest tests/test_multi_gpu.py tests/test_cpp_header.py -m gpu -x -q > gpurun_out/r02c_pytest_multi_n$N.log 2>&1; echo "pytest exit $?" >> gpurun_out/r02c_pytest_multi_n$N.log
tail -8 gpurun_out/r02c_pytest_multi_n$N.log
timeout 300 tests/cpp/_build/multi_survey 16777216 4096 0 nocheck > gpurun_out/r02c_multi_survey_cpp_n$N.txt 2>&1; echo "cpp exit $?" >> gpurun_out/r02c_multi_survey_cpp_n$N.txt
cat gpurun_out/r02c_multi_survey_cpp_n$N.txt
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 5 --warmup 3 > gpurun_out/r02c_bench_n$N.json 2> gpurun_out/r02c_bench_n$N.err; echo "bench exit $?"
tail -c 1500 gpurun_out/r02c_bench_n$N.err
cat gpurun_out/r02c_bench_n$N.json | head -c 600
