#!/bin/bash
(timeout 900 python -m pytest tests/test_preprocess_gpu.py tests/test_guard_zones_gpu.py tests/test_multi_gpu.py -m gpu -x -q 2>&1 | tail -3)
timeout 300 python scratch/cfg5_time.py 2>&1 | tail -10
timeout 300 python scratch/fuzz_pre.py 41 2000 | tail -3
