python -m pytest tests/test_step_gpu.py -m gpu -x -q 2>&1 | tail -3
for l in 1 2 4 8; do
  python bench.py --layout $l --steps 3 --warmup 3 --no-e2e --no-cpu --no-extras | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print('layout',d['step_layout'],'Gcell/s',round(d['value'],1),'ms',round(d['ms_per_step'],3),'checksum',d['checksum']['hash_xor'])"
done
