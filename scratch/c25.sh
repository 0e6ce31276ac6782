#!/bin/bash
(timeout 600 python -m pytest tests/test_cudashm_gpu.py tests/test_system_shm.py tests/test_http.py -m gpu -x -q 2>&1 | tail -5)
timeout 600 python bench.py --steps 100 --warmup 10 --no-cpu --no-e2e --also region_copy,cfg2_serialize,bf16 2>/dev/null | python -c "
import sys,json; d=json.loads(sys.stdin.read()); a=d['also']
print('region', json.dumps(a['region_copy'])[:1400])
print('cfg2', a['cfg2_serialize']['us_per_launch'], a['cfg2_serialize']['roofline']['frac'], a['cfg2_serialize']['us_per_call_isolated'], a['cfg2_serialize']['at_16M_elements'], a['cfg2_serialize']['bit_exact_vs_oracle_rank0'])
print('bf16', a['bf16']['serialize_us'], a['bf16']['deserialize_us'], a['bf16']['roofline']['frac'], a['bf16']['roofline_deserialize']['frac'], a['bf16']['bit_exact_vs_oracle_rank0'])"
