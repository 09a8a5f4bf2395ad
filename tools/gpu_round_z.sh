set -x
python -m pytest tests -m gpu -x -q 2>&1 | tail -4 | tee gpurun_out/r01z_pytest_gpu.log
python __graft_entry__.py --smoke 2>&1 | tail -2 | tee gpurun_out/r01z_smoke.log
python bench.py > gpurun_out/r01z_bench.json 2> gpurun_out/r01z_bench.err; echo bench rc=$?
tail -c 300 gpurun_out/r01z_bench.err
python tools/step_sweep.py | head -12 | tee gpurun_out/r01z_step_sweep_auto.txt
