#!/bin/bash
(timeout 900 python -m pytest tests/test_bytes_gpu.py tests/test_http.py tests/test_cudashm_gpu.py tests/test_postprocess.py -m gpu -x -q 2>&1 | tail -4)
timeout 300 python scratch/small_lat.py 2>&1 | tail -6
