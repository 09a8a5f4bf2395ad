set -x
ncu --set full --clock-control none --import-source on -k regex:"unknown_step_kernel" -s 3 -c 1 -o gpurun_out/r01m_unknown_step -f python bench.py --steps 1 --warmup 3 --no-cpu --no-e2e > gpurun_out/ncu_full_m2.log 2>&1; echo ncu rc=$?
python bench.py > gpurun_out/r01m_bench.json 2> gpurun_out/r01m_bench.err; echo bench rc=$?; wc -l gpurun_out/r01m_bench.json
