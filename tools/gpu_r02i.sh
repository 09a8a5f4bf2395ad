#!/bin/bash
set -x
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/r02i_pytest_gpu.log 2>&1; echo "pytest exit $?" >> gpurun_out/r02i_pytest_gpu.log
tail -5 gpurun_out/r02i_pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r02i_smoke.log 2>&1; echo "smoke exit $?" >> gpurun_out/r02i_smoke.log; cat gpurun_out/r02i_smoke.log
timeout 600 python bench.py > gpurun_out/r02i_bench.json 2> gpurun_out/r02i_bench.err; echo "bench exit $?"; tail -c 800 gpurun_out/r02i_bench.err
timeout 600 python bench.py --impl reference > gpurun_out/r02i_bench_reference_arm.json 2>> gpurun_out/r02i_bench.err; echo "ref arm exit $?"; cat gpurun_out/r02i_bench_reference_arm.json | head -c 400
