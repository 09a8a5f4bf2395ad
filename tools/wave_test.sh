# does wave quantisation explain the LPP-dependent utilisation?  whole numbers of waves per SM:
run() { python bench.py --lpp $1 --soups $2 --steps 3 --warmup 3 --no-e2e --no-cpu --no-extras | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print('lpp',d['lanes_per_pattern'],'soups',d['config']['soups_per_gpu'],'Gcell/s',round(d['value'],1),'util',round(d['roofline']['pipe_utilisation'],4),'ms',round(d['ms_per_step'],3))"; }
run 4 1022976    # 27 waves x 592 CTAs x 64 patterns
run 4 1060864    # 28 waves
run 8 1041920    # 44 waves x 740 CTAs x 32
run 16 1037184   # 73 waves x 888 CTAs x 16
run 16 1051392   # 74 waves
run 32 1051648   # 1184 CTAs/wave x 8 patterns x 111 waves
