set -x
python -m pytest tests -m gpu -x -q 2>&1 | tail -4 | tee gpurun_out/r01m_pytest_gpu.log
ncu --set full --clock-control none --import-source on -k regex:"stable_pass_kernel" -s 2 -c 1 -o gpurun_out/r01m_stable_pass -f python bench.py --steps 1 --warmup 3 --no-cpu --no-e2e > gpurun_out/ncu_full_m.log 2>&1; echo ncu rc=$?
tail -3 gpurun_out/ncu_full_m.log
