set -x
python -m pytest tests -m gpu -x -q 2>&1 | tail -3 | tee gpurun_out/r01o_pytest_gpu.log
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --impl reference --gpus 2 --steps 2 --warmup 1 > gpurun_out/r01o_bench_n2_ref.json 2> gpurun_out/r01o_n2_ref.err; echo ref rc=$?
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus 2 --steps 5 --warmup 3 > gpurun_out/r01o_bench_n2.json 2> gpurun_out/r01o_n2.err; echo n2 rc=$?
tail -c 400 gpurun_out/r01o_n2.err
python - <<'PY'
import json
d=json.load(open('gpurun_out/r01o_bench_n2.json'))
print('N=2 value',d['value'],'e2e',d['e2e'] and d['e2e']['value'],'gather',d.get('survey_gather'))
n=d['other_configs'].get('next_rows',{})
print({k:(v.get('hbm_gbs') or v.get('value')) for k,v in n.items()})
PY
ncu --set full --clock-control none --import-source on -k regex:"interaction_kernel" -s 1 -c 1 -o gpurun_out/r01o_interaction_kernel -f python bench.py --steps 1 --warmup 3 --no-cpu --no-e2e > gpurun_out/ncu_full_o.log 2>&1; echo ncu rc=$?
