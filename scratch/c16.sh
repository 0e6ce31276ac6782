#!/bin/bash
mkdir -p gpurun_out
(timeout 600 python -m pytest tests/test_preprocess_gpu.py tests/test_multi_gpu.py -m gpu -x -q > gpurun_out/c16_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/c16_pytest.log)
(TCB_MMA_SPLIT=1 timeout 600 python -m pytest tests/test_preprocess_gpu.py -m gpu -x -q > gpurun_out/c16_pytest_split.log 2>&1; echo "pytest rc=$?" >> gpurun_out/c16_pytest_split.log)
B="python bench.py --steps 50 --warmup 5 --no-cpu --no-e2e"
run() { echo "== $1" >> gpurun_out/c16_matrix.log; env $1 timeout 120 $B --also cfg5_mixed 2>> gpurun_out/c16_matrix.log | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(d['ms_per_step'], d['roofline']['frac'], d['config']['parity_spot_check_all_ranks'], d['also']['cfg5_mixed']['ms_per_batch'], d['also']['cfg5_mixed']['parity_one_image_per_size_rank0'])" >> gpurun_out/c16_matrix.log 2>&1; }
run "X=1"
run "X=2"
tail -3 gpurun_out/c16_pytest.log; tail -3 gpurun_out/c16_pytest_split.log; cat gpurun_out/c16_matrix.log
