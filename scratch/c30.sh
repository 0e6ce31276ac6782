#!/bin/bash
(timeout 900 python -m pytest tests/test_bytes_gpu.py tests/test_cudashm_gpu.py tests/test_http.py tests/test_system_shm.py tests/test_pyobj.py -m "gpu or not gpu" -x -q 2>&1 | tail -3)
timeout 600 python bench.py --steps 20 --warmup 3 --no-cpu --no-e2e --also cfg2_serialize 2>/dev/null | python -c "
import sys,json; d=json.loads(sys.stdin.read()); print(json.dumps(d['also']['bytes_api_level'])[:700])"
