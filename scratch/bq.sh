#!/bin/bash
# quick GPU check: bytes tests + the three bench numbers
timeout 600 python -m pytest tests/test_bytes_gpu.py tests/test_cudashm_gpu.py tests/test_zz_no_alloc.py -x -q -m gpu 2>&1 | tail -4
timeout 200 python bench.py --no-cpu --steps 10 --warmup 3 2>/dev/null | python -c "
import sys,json
for l in sys.stdin:
    if l.startswith('{'):
        d=json.loads(l); print('fused ms', d['ms_per_step'], 'ser us', d['also']['cfg2_serialize']['us_per_launch'], 'des ms', d['also']['cfg4_deserialize']['ms_per_call'])
"
