#!/bin/bash
# round 2, first GPU pass: parity suite, short bench, reference arm, launch list, ncu captures of the secondary kernels
set -x
mkdir -p gpurun_out
nvidia-smi -L > gpurun_out/r02a_gpus.txt
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/r02a_pytest_gpu.log 2>&1; echo "pytest exit $?" >> gpurun_out/r02a_pytest_gpu.log
tail -15 gpurun_out/r02a_pytest_gpu.log
timeout 600 python bench.py --steps 5 --warmup 3 > gpurun_out/r02a_bench.json 2> gpurun_out/r02a_bench.err; echo "bench exit $?"
tail -c 3000 gpurun_out/r02a_bench.err
timeout 300 python tools/ncu_targets.py > gpurun_out/r02a_targets.log 2>&1; echo "targets exit $?"; cat gpurun_out/r02a_targets.log
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"enumerate_kernel_x|weld_step_kernel|step_match_kernel|unknown_step_kernel" -c 6 -o gpurun_out/r02a_secondary python tools/ncu_targets.py > gpurun_out/r02a_ncu.log 2>&1; echo "ncu exit $?"
tail -5 gpurun_out/r02a_ncu.log
