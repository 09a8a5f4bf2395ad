#!/bin/bash
N=${1:-2}
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_multi_gpu.py tests/test_cpp_header.py -m gpu -x -q 2>&1 | tail -3
export LIFEB200_TRACE_GATHER=1
timeout 300 tests/cpp/_build/multi_survey 16777216 4096 0 nocheck 2>&1 | tee gpurun_out/r02e_cpp_trace_n$N.txt
unset LIFEB200_TRACE_GATHER
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29513 bench.py --gpus $N --steps 5 --warmup 3 > gpurun_out/r02e_bench_n$N.json 2> gpurun_out/r02e_bench_n$N.err; echo "bench exit $?"
python -c "
import json; d=json.load(open('gpurun_out/r02e_bench_n$N.json')); print(d['value'], d['e2e']['value'], d['parity']['equal']); print(d['survey_gather'])"
