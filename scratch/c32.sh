#!/bin/bash
(timeout 900 python -m pytest tests/test_bytes_gpu.py tests/test_zz_no_alloc.py tests/test_guard_zones_gpu.py tests/test_multi_gpu.py -m gpu -x -q 2>&1 | tail -3)
for e in 0 1 0 1; do
TCB_SER_HALF_TILES=$e timeout 300 python bench.py --steps 30 --warmup 5 --no-cpu --no-e2e --also cfg2_serialize 2>/dev/null | python -c "
import sys,json; d=json.loads(sys.stdin.read()); a=d['also']['cfg2_serialize']; print($e, a['us_per_launch'], a['roofline']['frac'], a['us_per_call_isolated'], a['bit_exact_vs_oracle_rank0'], a['at_16M_elements']['roofline']['frac'])"
done
timeout 300 python scratch/fuzz_bytes.py 31 600 | tail -3
TCB_SER_HALF_TILES=1 timeout 300 python scratch/fuzz_bytes.py 32 600 | tail -3
