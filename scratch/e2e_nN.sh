#!/bin/bash
N=$1
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29519 bench.py --gpus $N --no-cpu --also none 2>/dev/null | grep '^{' | python -c "
import sys,json; d=json.loads(sys.stdin.read()); e=d['e2e']
print(d['n_gpus'], d['value'], 'e2e', e['value'], 'h2d ceil', e['h2d_ceiling_gbs'], e['frac_of_h2d_ceiling'], 'bidir ceil', e['bidir_ceiling_gbs'], e['frac_of_bidir_ceiling'], 'pageable', e['from_pageable'].get('value'), e['affinity'])"
