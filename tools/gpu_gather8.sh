python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29578 bench.py --gpus 8 --steps 5 --warmup 3 2>/dev/null > gpurun_out/r01z4_bench_n8.json
python -c "
import json
d=json.load(open('gpurun_out/r01z4_bench_n8.json')); print(8,'value',round(d['value']),'e2e',round(d['e2e']['value']),'gather',d['survey_gather'])"
