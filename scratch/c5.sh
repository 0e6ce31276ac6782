#!/bin/bash
mkdir -p gpurun_out
(timeout 600 python -m pytest tests/test_preprocess_gpu.py tests/test_multi_gpu.py tests/test_cudashm_gpu.py -m gpu -x -q > gpurun_out/c5_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/c5_pytest.log)
B="python bench.py --steps 30 --warmup 5 --no-cpu --no-e2e"
for per in ""; do
  echo "== MMA=1 CTAS_PER_SM=$per" >> gpurun_out/c5_matrix.log
  TCB_DEBUG=1 TCB_FUSED_CTAS_PER_SM=$per timeout 120 $B --also cfg5_mixed 2>> gpurun_out/c5_matrix.log | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(d['ms_per_step'], d['roofline']['frac'], d['config']['parity_spot_check_all_ranks'], d['also'])" >> gpurun_out/c5_matrix.log 2>&1
done
B2="python bench.py --steps 3 --warmup 3 --no-cpu --no-e2e --also none"
$B2 > gpurun_out/c5_plain_mma.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:k_preprocess_fused -s 3 -c 1 -o gpurun_out/c5_mma $B2 > gpurun_out/c5_ncu_mma.log 2>&1
tail -5 gpurun_out/c5_pytest.log; cat gpurun_out/c5_matrix.log
