timeout 300 python -m pytest tests -m gpu -x -q -k "crop or golden or reference_api" 2>&1 | tail -n 15
for c in 2 3 4; do timeout 200 python bench.py --config $c --steps 20 --warmup 3 --no-cpu --no-e2e 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); k=d['kernels']
print('cfg',d['config']['cfg'],'step',round(d['ms_per_step'],3),{n:(round(v['ms'],3),round(v['frac'],3)) for n,v in k.items() if n in('crop','encode','decode')})"; done
