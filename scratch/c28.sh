#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -X faulthandler -m pytest tests/test_preprocess_gpu.py -m gpu -x -q > gpurun_out/c28_pytest.log 2>&1
echo "rc=$?" >> gpurun_out/c28_pytest.log
grep -n "Fatal\|File \|passed\|failed\|rc=" gpurun_out/c28_pytest.log | head -40
