#!/bin/bash
(timeout 600 python -m pytest tests/test_preprocess_gpu.py tests/test_cudashm_gpu.py -m gpu -x -q 2>&1 | tail -3)
for p in 4294967296 0; do
TCB_HOST_POOL_BYTES=$p timeout 300 python bench.py --steps 50 --warmup 5 --no-cpu --no-e2e --also cfg1_latency 2>/dev/null | python -c "
import sys,json; d=json.loads(sys.stdin.read()); a=d['also']['cfg1_latency']; print($p, a['value'], a['p10_us'], a['p90_us'], a['golden_sha256_prefix_matches'])"
done
