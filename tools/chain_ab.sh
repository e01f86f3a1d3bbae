#!/bin/bash
# A/B of wavefront chain count x persistent-grid cap (RT2_CHAINS, RT2_RAYS_PER_LANE) on the bench workload
for cfg in ${CONFIGS:-"2 0" "2 12" "2 15" "2 20" "3 12" "3 20" "4 20" "4 30" "2 0"}; do set -- $cfg
RT2_CHAINS=$1 RT2_RAYS_PER_LANE=$2 timeout 300 python bench.py --steps 8 --warmup 3 --no-cpu-baseline 2>&1 | tail -1 | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print('  chains $1 rays/lane $2:', round(d['value'],1), 'Mrays/s e2e', round(d['e2e']['value'],1))"; done
