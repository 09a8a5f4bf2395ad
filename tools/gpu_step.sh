python -m pytest tests/test_step_gpu.py tests/test_match_gpu.py -m gpu -x -q 2>&1 | tail -2
python tools/step_sweep.py | tee gpurun_out/r01z_step_sweep.txt
