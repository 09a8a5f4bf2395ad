python -m pytest tests -m gpu -x -q 2>&1 | tail -3 | tee gpurun_out/r01z3_pytest_gpu.log
python __graft_entry__.py --smoke 2>&1 | tail -1
python tools/multi_thresholds.py | tee gpurun_out/r01z3_multi_thresholds.txt
python -m torch.distributed.run --nnodes=1 --nproc-per-node 1 --master-addr 127.0.0.1 --master-port 29561 bench.py --gpus 1 --steps 3 --warmup 3 --no-cpu 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print('value',round(d['value']),'c4',d['other_configs']['config4']['placements_per_s'],'c5',d['other_configs']['config5']['value'],'gather',d.get('survey_gather'))"
