#!/bin/bash
# round 2, second GPU pass (1 GPU): parity suite on the new kernels, bench, sanitizer runs of the smoke workload,
# launch list of the bench, full captures of the step kernel (roofline.traffic source) and the new layout-8 kernels
set -x
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/r02b_pytest_gpu.log 2>&1; echo "pytest exit $?" >> gpurun_out/r02b_pytest_gpu.log
tail -8 gpurun_out/r02b_pytest_gpu.log
timeout 600 python bench.py --steps 5 --warmup 3 > gpurun_out/r02b_bench.json 2> gpurun_out/r02b_bench.err; echo "bench exit $?"
tail -c 1500 gpurun_out/r02b_bench.err
timeout 600 compute-sanitizer --tool memcheck --error-exitcode 7 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r02b_sanitizer_memcheck.txt 2>&1; echo "memcheck exit $?" >> gpurun_out/r02b_sanitizer_memcheck.txt
timeout 900 compute-sanitizer --tool racecheck --error-exitcode 7 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r02b_sanitizer_racecheck.txt 2>&1; echo "racecheck exit $?" >> gpurun_out/r02b_sanitizer_racecheck.txt
tail -n 4 gpurun_out/r02b_sanitizer_memcheck.txt; tail -n 4 gpurun_out/r02b_sanitizer_racecheck.txt
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r02b_launches_bench_steps3.csv python bench.py --steps 3 --warmup 3 --no-cpu > gpurun_out/r02b_ncu_bench.log 2>&1; echo "ncu launches exit $?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"step_kernel_x" -c 1 -s 4 -o gpurun_out/r02b_step_kernel_x8 python bench.py --steps 2 --warmup 3 --no-cpu --no-extras --no-e2e --no-survey > gpurun_out/r02b_ncu_step.log 2>&1; echo "ncu step exit $?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"weld_step_kernel_x|step_match_kernel_x|enumerate_kernel_x" -c 4 -o gpurun_out/r02b_secondary python tools/ncu_targets.py enumerate weld match > gpurun_out/r02b_ncu_secondary.log 2>&1; echo "ncu secondary exit $?"
tail -3 gpurun_out/r02b_ncu_secondary.log
