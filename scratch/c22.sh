#!/bin/bash
for e in 0 1 2 3 4 6 0 1 2 3 4 6; do
TCB_FUSED_L2PF=$e python bench.py --steps 100 --warmup 10 --no-cpu --no-e2e --also none 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.read()); print($e, d['ms_per_step'], d['roofline']['frac'], d['config']['parity_spot_check_all_ranks'])"
done
