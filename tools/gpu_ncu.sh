# ncu evidence for the bench command (run only after the same command exited 0 without ncu)
python bench.py --steps 3 --warmup 3 --no-cpu --no-e2e > gpurun_out/bench_plain.json 2> gpurun_out/bench_plain.err; echo plain rc=$?
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r01j_launches.csv python bench.py --steps 3 --warmup 3 --no-cpu --no-e2e > gpurun_out/ncu_launch.log 2>&1; echo ncu1 rc=$?
ncu --set full --clock-control none --import-source on -k regex:step_kernel_x -s 3 -c 1 -o gpurun_out/r01j_step_kernel_x8 -f python bench.py --steps 2 --warmup 3 --no-cpu --no-e2e --no-extras > gpurun_out/ncu_full.log 2>&1; echo ncu2 rc=$?
ncu --set full --clock-control none --import-source on -k regex:survey_kernel_x -c 1 -o gpurun_out/r01j_survey_kernel_x8 -f python bench.py --steps 1 --warmup 3 --no-cpu --no-e2e > gpurun_out/ncu_full2.log 2>&1; echo ncu3 rc=$?
