#!/bin/bash
# round 2, final tree, 2 GPUs: multi-GPU tests, the C++ program, the bench under torchrun (both arms) and single-process
set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_multi_gpu.py tests/test_cpp_header.py -m gpu -x -q > gpurun_out/r02v_pytest_multi_n2.log 2>&1; echo "pytest exit $?" >> gpurun_out/r02v_pytest_multi_n2.log
tail -4 gpurun_out/r02v_pytest_multi_n2.log
timeout 300 tests/cpp/_build/multi_survey 16777216 4096 2 nocheck > gpurun_out/r02v_multi_survey_cpp_n2.txt 2>&1; echo "cpp exit $?" >> gpurun_out/r02v_multi_survey_cpp_n2.txt
cat gpurun_out/r02v_multi_survey_cpp_n2.txt
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 bench.py --impl reference --gpus 2 --steps 3 --warmup 1 > gpurun_out/r02v_bench_reference_arm_n2.json 2> gpurun_out/r02v_bench_reference_arm_n2.err; echo "ref arm n2 exit $?"; cat gpurun_out/r02v_bench_reference_arm_n2.json | cut -c1-300
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29513 bench.py --gpus 2 --steps 5 --warmup 3 > gpurun_out/r02v_bench_n2.json 2> gpurun_out/r02v_bench_n2.err; echo "bench n2 exit $?"
python -c "
import json; d=json.load(open('gpurun_out/r02v_bench_n2.json')); print(d['n_gpus'], d['value'], d['e2e']['value'], d['e2e']['ms_per_step_by_rank'], d['parity']['equal']); print(d['survey_gather'])"
timeout 600 python bench.py --gpus 2 --steps 5 --warmup 3 --no-extras > gpurun_out/r02v_bench_single_process_n2.json 2> gpurun_out/r02v_bench_single_process_n2.err; echo "single-process exit $?"
python -c "
import json; d=json.load(open('gpurun_out/r02v_bench_single_process_n2.json')); print(d['n_gpus'], d['value'], d['e2e']['value'], d['parity'])"
