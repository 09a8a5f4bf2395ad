# LPP sweep of the step kernel (config 2), device-timed; prints value / pipe_utilisation per lanes-per-pattern
for l in 1 2 4 8 16; do
  python bench.py --lpp $l --steps 3 --warmup 3 --no-e2e --no-cpu --no-extras | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print('lpp',d['lanes_per_pattern'],'Gcell/s',round(d['value'],1),'util',round(d['roofline']['pipe_utilisation'],4),'ms',round(d['ms_per_step'],3))"
done
