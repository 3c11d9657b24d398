#!/bin/bash
for w in c5 c4; do for spp in 1 2 4 8; do
 timeout 300 python bench.py --workload $w --no-cpu --steps 2 --warmup 2 --samples-per-pass $spp 2>/dev/null | python -c "
import json,sys;d=json.loads(sys.stdin.read());print('$w spp/pass $spp %.0f Mrays/s %.0f ms e2e %.0f'%(d['value'],d['ms_per_step'],d['e2e']['value']),d['stage_ms'])"
done; done
