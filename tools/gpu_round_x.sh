set -x
python -m pytest tests -m gpu -x -q 2>&1 | tail -4 | tee gpurun_out/r01x_pytest_gpu.log
python __graft_entry__.py --smoke 2>&1 | tail -2 | tee gpurun_out/r01x_smoke.log
python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/r01x_bench_reference_arm.json 2> gpurun_out/r01x_ref.err; echo ref rc=$?
python bench.py > gpurun_out/r01x_bench.json 2> gpurun_out/r01x_bench.err; echo bench rc=$?
tail -c 300 gpurun_out/r01x_bench.err
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r01x_launches_bench_steps3.csv python bench.py --steps 3 --warmup 3 --no-cpu > gpurun_out/ncu_x.log 2>&1; echo ncu rc=$?
ncu --set full --clock-control none --import-source on -k regex:"^unknown_step_kernel" -s 2 -c 1 -o gpurun_out/r01x_unknown_step -f python bench.py --steps 1 --warmup 3 --no-cpu --no-e2e > gpurun_out/ncu_full_x1.log 2>&1; echo ncu rc=$?
