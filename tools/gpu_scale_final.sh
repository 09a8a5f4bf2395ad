set -x
python bench.py --gpus 1 --steps 5 --warmup 3 --no-extras > gpurun_out/r01z_scale_n1.json 2> gpurun_out/r01z_scale_n1.err; echo n1 rc=$?
for n in 2 4 8; do
python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 2954$n bench.py --gpus $n --steps 5 --warmup 3 > gpurun_out/r01z_scale_n$n.json 2> gpurun_out/r01z_scale_n$n.err; echo n$n rc=$?
done
python - <<'PY'
import json
for n in (1,2,4,8):
    d=json.load(open(f'gpurun_out/r01z_scale_n{n}.json'))
    print(n,'value',round(d['value']),'e2e',d['e2e'] and round(d['e2e']['value']),'ms',round(d['ms_per_step'],3),'by_rank',d['e2e'].get('ms_per_step_by_rank'))
PY
