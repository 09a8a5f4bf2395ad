python -m pytest tests -m gpu -x -q 2>&1 | tail -3
python bench.py --no-cpu --no-extras | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print('value',round(d['value'],1),'e2e',round(d['e2e']['value'],1),'ms',round(d['ms_per_step'],3),'util',round(d['roofline']['pipe_utilisation'],4),'checksum',d['checksum'])"
