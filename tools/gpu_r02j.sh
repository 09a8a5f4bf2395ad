#!/bin/bash
set -x
N=${1:-2}
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_match_gpu.py tests/test_host_calls_gpu.py -m gpu -x -q 2>&1 | tail -4
timeout 600 python bench.py --gpus $N --steps 5 --warmup 3 > gpurun_out/r02j_bench_single_process_n$N.json 2> gpurun_out/r02j_bench_single_process_n$N.err; echo "bench exit $?"; tail -c 1500 gpurun_out/r02j_bench_single_process_n$N.err
python -c "
import json; d=json.load(open('gpurun_out/r02j_bench_single_process_n$N.json')); print(d['n_gpus'], d['value'], d['e2e'], d['parity']); print(d['survey_gather'])"
