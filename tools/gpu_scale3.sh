set -x
python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29518 bench.py --gpus 8 --steps 5 --warmup 3 > gpurun_out/r01r_bench_n8.json 2> gpurun_out/r01r_n8.err; echo n8 rc=$?
python - <<'PY'
import json
d=json.load(open('gpurun_out/r01r_bench_n8.json'))
print('value',round(d['value']),'e2e',d['e2e'])
PY
