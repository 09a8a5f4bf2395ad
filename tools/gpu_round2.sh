python -m pytest tests -m gpu -x -q 2>&1 | tail -3
python bench.py > gpurun_out/bench_r01h.json 2> gpurun_out/bench_r01h.err; echo bench rc=$?
bash tools/lpp_sweep.sh
./tools/ubench_step > gpurun_out/r01h_ubench_step.txt 2>&1
python -c "
import json
d=json.load(open('gpurun_out/bench_r01h.json'))
print('value',d['value'],'e2e',d['e2e']['value'],'util',d['roofline']['pipe_utilisation'],'frac',d['roofline']['frac'])
for k,v in d['other_configs'].items(): print(k, json.dumps(v)[:400])
"
