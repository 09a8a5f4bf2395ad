set -x
python -m pytest tests -m gpu -x -q 2>&1 | tail -4
python bench.py > gpurun_out/bench_r01g.json 2> gpurun_out/bench_r01g.err; echo bench rc=$?
python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_r01g_ref.json 2>> gpurun_out/bench_r01g.err; echo ref rc=$?
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r01g_launches.csv python bench.py --steps 3 --warmup 3 --no-cpu --no-e2e > gpurun_out/ncu_launch.log 2>&1; echo ncu1 rc=$?
ncu --set full --clock-control none --import-source on -k regex:step_kernel -s 3 -c 1 -o gpurun_out/r01g_step_kernel -f python bench.py --steps 2 --warmup 3 --no-cpu --no-e2e --no-extras > gpurun_out/ncu_full.log 2>&1; echo ncu2 rc=$?
ncu --set full --clock-control none --import-source on -k regex:"match_kernel|interaction_offsets_kernel" -c 2 -o gpurun_out/r01g_locate_kernels -f python bench.py --steps 1 --warmup 3 --no-cpu --no-e2e > gpurun_out/ncu_full2.log 2>&1; echo ncu3 rc=$?
cat gpurun_out/bench_r01g.json
