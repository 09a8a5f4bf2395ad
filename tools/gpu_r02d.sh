#!/bin/bash
set -x
N=${1:-2}
mkdir -p gpurun_out
timeout 600 python tools/multi_probe.py > gpurun_out/r02d_probe_single_n$N.txt 2>&1; echo "exit $?"; cat gpurun_out/r02d_probe_single_n$N.txt
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29512 tools/multi_probe.py > gpurun_out/r02d_probe_ranks_n$N.txt 2>&1; echo "exit $?"; grep -v "^\*\|OMP" gpurun_out/r02d_probe_ranks_n$N.txt
timeout 600 python -m pytest tests/test_multi_gpu.py -m gpu -x -q 2>&1 | tail -5
timeout 300 tests/cpp/_build/multi_survey 16777216 4096 0 nocheck
