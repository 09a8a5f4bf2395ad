#!/bin/bash
# round 2, evidence pass on the tree with the 11-instruction pack and the re-measured layout thresholds
set -x
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/r02v_pytest_gpu.log 2>&1; echo "pytest exit $?" >> gpurun_out/r02v_pytest_gpu.log
tail -5 gpurun_out/r02v_pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r02v_smoke.log 2>&1; echo "smoke exit $?" >> gpurun_out/r02v_smoke.log; cat gpurun_out/r02v_smoke.log
timeout 600 python bench.py > gpurun_out/r02v_bench.json 2> gpurun_out/r02v_bench.err; echo "bench exit $?"; tail -c 800 gpurun_out/r02v_bench.err
timeout 600 python bench.py --impl reference > gpurun_out/r02v_bench_reference_arm.json 2>> gpurun_out/r02v_bench.err; echo "ref arm exit $?"
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r02v_launches_bench_steps3.csv python bench.py --steps 3 --warmup 3 --no-cpu > gpurun_out/r02v_ncu_bench.log 2>&1; echo "ncu launches exit $?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"step_kernel_x" -c 1 -s 4 -o gpurun_out/r02v_step_kernel_x8 python bench.py --steps 2 --warmup 3 --no-cpu --no-extras --no-e2e --no-survey > gpurun_out/r02v_ncu_step.log 2>&1; echo "ncu step exit $?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"step_match_kernel_x|enumerate_kernel_x" -c 2 -o gpurun_out/r02v_secondary python tools/ncu_targets.py enumerate match > gpurun_out/r02v_ncu_secondary.log 2>&1; echo "ncu secondary exit $?"
timeout 900 ncu --set full --clock-control none --replay-mode application -k regex:"weld_step_kernel_x" -c 1 -o gpurun_out/r02v_weld python tools/ncu_targets.py weld > gpurun_out/r02v_ncu_weld.log 2>&1; echo "ncu weld exit $?"
tail -3 gpurun_out/r02v_ncu_weld.log
