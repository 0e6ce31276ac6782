#!/bin/bash
# staging threads sweep: region set/get from pageable memory and the pageable e2e
for t in 6 10 14; do
echo "== TCB_STAGING_THREADS=$t"
TCB_STAGING_THREADS=$t python bench.py --steps 20 --warmup 3 --no-cpu --also region_copy 2>/dev/null | python -c "
import sys,json; d=json.loads(sys.stdin.read()); print(d['e2e']['value'], d['e2e']['from_pageable']['value'], d['also']['region_copy']['staged'])"
done
