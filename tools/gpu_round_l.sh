set -x
python -m pytest tests -m gpu -x -q 2>&1 | tail -4 | tee gpurun_out/r01l_pytest_gpu.log
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r01l_smoke.log 2>&1; echo smoke rc=$?
python bench.py --impl reference > gpurun_out/r01l_bench_reference_arm.json 2> gpurun_out/r01l_bench_ref.err; echo ref rc=$?
python bench.py > gpurun_out/r01l_bench.json 2> gpurun_out/r01l_bench.err; echo bench rc=$?
python bench.py --steps 3 --warmup 3 --no-cpu --no-e2e > gpurun_out/bench_plain.json 2> gpurun_out/bench_plain.err; echo plain rc=$?
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r01l_launches_bench_steps3.csv python bench.py --steps 3 --warmup 3 --no-cpu --no-e2e > gpurun_out/ncu_launch.log 2>&1; echo ncu1 rc=$?
cat gpurun_out/r01l_bench.json
