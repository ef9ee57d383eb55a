#!/bin/bash
# A/B of the sorted fast path of the update on the N=1 bench
for v in 1 0 1 0; do PER_SORTED_UPDATE=$v timeout 300 python bench.py --steps 2000 --warmup 20 --no-cpu-baseline 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1])
print('PER_SORTED_UPDATE=$v', 'value %.3f M/s' % (d['value']/1e6), 'us/step %.2f' % (d['ms_per_step']*1e3), 'eager %.3f' % ((d.get('value_eager_launches') or 0)/1e6))"; done
