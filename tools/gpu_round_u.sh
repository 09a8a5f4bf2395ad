set -x
python -m pytest tests -m gpu -x -q 2>&1 | tail -4 | tee gpurun_out/r01u_pytest_gpu.log
python __graft_entry__.py --smoke 2>&1 | tail -2 | tee gpurun_out/r01u_smoke.log
python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/r01u_bench_reference_arm.json 2> gpurun_out/r01u_ref.err; echo ref rc=$?
python bench.py > gpurun_out/r01u_bench.json 2> gpurun_out/r01u_bench.err; echo bench rc=$?
tail -c 300 gpurun_out/r01u_bench.err
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r01u_launches_bench_steps3.csv python bench.py --steps 3 --warmup 3 --no-cpu > gpurun_out/ncu_u.log 2>&1; echo ncu rc=$?
ncu --set full --clock-control none --import-source on -k regex:"^stable_pass_kernel" -s 2 -c 1 -o gpurun_out/r01u_stable_pass -f python bench.py --steps 1 --warmup 3 --no-cpu --no-e2e > gpurun_out/ncu_full_u1.log 2>&1; echo ncu rc=$?
ncu --set full --clock-control none --import-source on -k regex:"^match_kernel" -s 1 -c 1 -o gpurun_out/r01u_match_kernel -f python bench.py --steps 1 --warmup 3 --no-cpu --no-e2e > gpurun_out/ncu_full_u2.log 2>&1; echo ncu rc=$?
ncu --set full --clock-control none --import-source on -k regex:"^interaction_offsets_kernel" -s 1 -c 1 -o gpurun_out/r01u_interaction_offsets -f python bench.py --steps 1 --warmup 3 --no-cpu --no-e2e > gpurun_out/ncu_full_u3.log 2>&1; echo ncu rc=$?
