#!/bin/bash
# round 2, final tree, 8-GPU pass: multi-GPU tests, the C++ program on 8 / 4 GPUs, the bench under torchrun at N = 8 and 4
set -x
mkdir -p gpurun_out
nvidia-smi -L > gpurun_out/r02v8_gpus.txt
timeout 900 python -m pytest tests/test_multi_gpu.py tests/test_cpp_header.py -m gpu -x -q > gpurun_out/r02v8_pytest_multi_n8.log 2>&1; echo "pytest exit $?" >> gpurun_out/r02v8_pytest_multi_n8.log
tail -4 gpurun_out/r02v8_pytest_multi_n8.log
for g in 8 4; do
  LIFEB200_TRACE_GATHER=1 timeout 300 tests/cpp/_build/multi_survey 16777216 4096 $g nocheck > gpurun_out/r02v8_multi_survey_cpp_n$g.txt 2>&1; echo "cpp exit $?" >> gpurun_out/r02v8_multi_survey_cpp_n$g.txt
  cat gpurun_out/r02v8_multi_survey_cpp_n$g.txt
done
for g in 8 4; do
  timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $g --master-addr 127.0.0.1 --master-port 2951$g bench.py --gpus $g --steps 5 --warmup 3 > gpurun_out/r02v8_bench_n$g.json 2> gpurun_out/r02v8_bench_n$g.err; echo "bench n$g exit $?"
  python -c "
import json; d=json.load(open('gpurun_out/r02v8_bench_n$g.json')); print(d['n_gpus'], d['value'], d['e2e']['value'], d['e2e']['ms_per_step_by_rank'], d['parity']['equal']); print(d['survey_gather'])"
done
