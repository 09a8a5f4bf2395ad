set -x
for n in 8 4; do
python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 2951$n bench.py --gpus $n --steps 5 --warmup 3 > gpurun_out/r01p_bench_n$n.json 2> gpurun_out/r01p_n$n.err; echo n$n rc=$?
done
python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29520 bench.py --impl reference --gpus 8 --steps 2 --warmup 1 > gpurun_out/r01p_bench_n8_ref.json 2> gpurun_out/r01p_n8_ref.err; echo ref rc=$?
python - <<'PY'
import json
for n in (8,4):
    d=json.load(open(f'gpurun_out/r01p_bench_n{n}.json'))
    print(n,'value',d['value'],'e2e',d['e2e'] and d['e2e']['value'],'ms',d['ms_per_step'],'clocks',d['clocks'],'gather',d.get('survey_gather'))
print(open('gpurun_out/r01p_bench_n8_ref.json').read()[:300])
PY
