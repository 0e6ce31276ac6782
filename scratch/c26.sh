#!/bin/bash
(timeout 600 python -m pytest tests/test_cudashm_gpu.py tests/test_preprocess_gpu.py tests/test_system_shm.py -m gpu -x -q 2>&1 | tail -4)
timeout 300 python scratch/pg_probe.py 2>&1 | tail -10
