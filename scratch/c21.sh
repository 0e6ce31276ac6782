#!/bin/bash
for k in 50 100 200 300 500 1000; do
python bench.py --steps $k --warmup 10 --no-cpu --no-e2e --also none 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.read()); print($k, d['ms_per_step'], d['roofline']['frac'], d['clocks'])"
done
