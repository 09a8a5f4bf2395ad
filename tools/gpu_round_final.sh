set -x
python -m pytest tests -m gpu -x -q 2>&1 | tail -3 | tee gpurun_out/r01_final_pytest_gpu.log
python __graft_entry__.py --smoke 2>&1 | tail -1 | tee gpurun_out/r01_final_smoke.log
python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/r01_final_bench_reference_arm.json 2>/dev/null; echo ref rc=$?
python bench.py > gpurun_out/r01_final_bench.json 2> gpurun_out/r01_final_bench.err; echo bench rc=$?
tail -c 300 gpurun_out/r01_final_bench.err
