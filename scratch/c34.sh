#!/bin/bash
(timeout 900 python -m pytest tests/test_bytes_gpu.py tests/test_cudashm_gpu.py tests/test_guard_zones_gpu.py -m gpu -x -q 2>&1 | tail -3)
timeout 300 python scratch/fuzz_bytes.py 51 800 | tail -3
TCB_TIMING=1 timeout 300 python bench.py --steps 5 --warmup 3 --no-cpu --no-e2e --also cfg4_deserialize,long_elements 2> gpurun_out/c34_err.log | python -c "
import sys,json; d=json.loads(sys.stdin.read()); a=d['also']['cfg4_deserialize']; print(a['ms_per_call'], a['roofline']['frac'], a['decoded_equals_input']); print({k:(v['ms_per_call'],v['decoded_equals_input']) for k,v in d['also']['long_elements'].items() if isinstance(v,dict) and 'ms_per_call' in v})"
grep "deserialize" gpurun_out/c34_err.log | sort | uniq -c | sort -rn | head -12
