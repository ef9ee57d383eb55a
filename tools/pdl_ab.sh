#!/bin/bash
# A/B of programmatic dependent launch on the N=1 bench (PER_PDL=0 switches it off in bench.py)
for v in ${PDL_SET:-7 0}; do PER_PDL=$v timeout 300 python bench.py --steps 2000 --warmup 20 --no-cpu-baseline 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1])
print('PER_PDL=$v', 'value %.3f M/s' % (d['value']/1e6), 'us/step %.2f' % (d['ms_per_step']*1e3), 'gather us %.2f' % d['roofline']['avg_launch_us'], 'frac %.3f' % d['roofline']['frac'], 'eager %.3f' % ((d.get('value_eager_launches') or 0)/1e6))"; done
