set -x
python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/r01z_bench_reference_arm.json 2> gpurun_out/r01z_ref.err; echo ref rc=$?
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r01z_launches_bench_steps3.csv python bench.py --steps 3 --warmup 3 --no-cpu > gpurun_out/ncu_z.log 2>&1; echo ncu rc=$?
ncu --set full --clock-control none --import-source on -k regex:"^step_kernel_x" -s 3 -c 1 -o gpurun_out/r01z_step_kernel_x8 -f python bench.py --steps 1 --warmup 3 --no-cpu --no-e2e --no-extras > gpurun_out/ncu_full_z.log 2>&1; echo ncu rc=$?
