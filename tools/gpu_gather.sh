for n in 2; do
python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 2957$n bench.py --gpus $n --steps 3 --warmup 3 2>/dev/null > gpurun_out/r01z4_bench_n$n.json
python -c "
import json
d=json.load(open('gpurun_out/r01z4_bench_n$n.json')); print($n,'value',round(d['value']),'e2e',round(d['e2e']['value']),'gather',d['survey_gather'])"
done
