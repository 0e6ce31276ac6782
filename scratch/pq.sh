#!/bin/bash
# quick GPU check: preprocess tests + fused bench number
timeout 900 python -m pytest tests/test_preprocess_gpu.py -x -q -m gpu 2>&1 | tail -4
timeout 200 python bench.py --no-cpu --no-also --steps 50 --warmup 5 2>/dev/null | python -c "
import sys,json
for l in sys.stdin:
    if l.startswith('{'):
        d=json.loads(l); print('fused ms', d['ms_per_step'], 'frac', d['roofline']['frac'])
"
