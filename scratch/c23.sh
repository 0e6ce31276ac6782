#!/bin/bash
(timeout 600 python -m pytest tests/test_preprocess_gpu.py tests/test_guard_zones_gpu.py -m gpu -x -q 2>&1 | tail -3)
for e in 0 1 0 1; do
TCB_FUSED_L2PF=$e python bench.py --steps 100 --warmup 10 --no-cpu --no-e2e --also cfg5_mixed 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.read()); print($e, d['ms_per_step'], d['roofline']['frac'], d['also']['cfg5_mixed']['ms_per_batch'], d['also']['cfg5_mixed']['roofline']['frac'], d['also']['cfg5_mixed']['parity_one_image_per_size_rank0'])"
done
for e in 0 1; do TCB_FUSED_L2PF=$e timeout 300 python scratch/cfg5_time.py 2>&1 | tail -10; done
