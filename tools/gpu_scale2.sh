set -x
nvidia-smi topo -m 2>&1 | head -14 > gpurun_out/r01q_topo.txt
lscpu | grep -E "NUMA|Socket|^CPU\(s\)|Model name" >> gpurun_out/r01q_topo.txt
python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29518 bench.py --gpus 8 --steps 5 --warmup 3 > gpurun_out/r01q_bench_n8.json 2> gpurun_out/r01q_n8.err; echo n8 rc=$?
LIFEB200_NO_NUMA_BIND=1 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29519 bench.py --gpus 8 --steps 5 --warmup 3 > gpurun_out/r01q_bench_n8_nobind.json 2> gpurun_out/r01q_n8_nobind.err; echo n8 rc=$?
python - <<'PY'
import json
for f in ('r01q_bench_n8','r01q_bench_n8_nobind'):
    d=json.load(open(f'gpurun_out/{f}.json'))
    print(f,'value',round(d['value']),'e2e',d['e2e'])
PY
cat gpurun_out/r01q_topo.txt
