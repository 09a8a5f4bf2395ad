python bench.py --lpp 32 --steps 3 --warmup 3 --no-e2e --no-cpu --no-extras | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print('lpp',d['lanes_per_pattern'],'Gcell/s',round(d['value'],1),'util',round(d['roofline']['pipe_utilisation'],4),'ms',round(d['ms_per_step'],3))"
for l in 8 16; do
ncu --set full --clock-control none --import-source on -k regex:step_kernel -s 3 -c 1 -o gpurun_out/r01h_step_lpp$l -f python bench.py --lpp $l --steps 2 --warmup 3 --no-cpu --no-e2e --no-extras > gpurun_out/ncu_lpp$l.log 2>&1; echo ncu lpp$l rc=$?
done
