python -m pytest tests/test_circuits_gpu.py -m gpu -x -q 2>&1 | tail -2
python tools/unknown_sweep.py 2>&1 | tee gpurun_out/r01w_unknown_sweep.txt
