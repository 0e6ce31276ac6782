#!/bin/bash
# compute-sanitizer over a reduced GPU test subset (the full-size cases are left out: the tools slow
# kernels by one to two orders of magnitude). Usage: san.sh memcheck|racecheck|synccheck|initcheck
TOOL=${1:-memcheck}
mkdir -p gpurun_out
SEL='not full_size and not 2gib and not large_pageable and not large_mixed and not long_elements and not gbytes_device and not second_process and not two_devices and not large_tensor'
FILES="tests/test_preprocess_gpu.py tests/test_bytes_gpu.py tests/test_bf16_gpu.py tests/test_zz_no_alloc.py tests/test_postprocess.py tests/test_cudashm_gpu.py"
timeout 1500 compute-sanitizer --tool $TOOL --target-processes all --print-limit 30 \
  --log-file gpurun_out/san_${TOOL}_%p.log \
  python -m pytest $FILES -m gpu -q -k "$SEL" -p no:cacheprovider > gpurun_out/san_${TOOL}_pytest.log 2>&1
echo "rc=$?" >> gpurun_out/san_${TOOL}_pytest.log
tail -5 gpurun_out/san_${TOOL}_pytest.log
for f in gpurun_out/san_${TOOL}_*.log; do echo "== $f"; grep -c "=========" $f; grep "ERROR SUMMARY\|RACECHECK SUMMARY\|Error:\|Invalid\|hazard" $f | sort | uniq -c | sort -rn | head -12; done
