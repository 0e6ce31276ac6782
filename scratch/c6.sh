#!/bin/bash
mkdir -p gpurun_out
(TCB_DEBUG=1 timeout 600 python -m pytest tests/test_preprocess_gpu.py tests/test_multi_gpu.py -m gpu -x -q > gpurun_out/c6_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/c6_pytest.log)
B="python bench.py --steps 30 --warmup 5 --no-cpu --no-e2e"
for tpw in 1 2; do
  echo "== TPW=$tpw" >> gpurun_out/c6_matrix.log
  TCB_DEBUG=1 TCB_MMA_TPW=$tpw timeout 120 $B --also cfg5_mixed 2>> gpurun_out/c6_matrix.log | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(d['ms_per_step'], d['roofline']['frac'], d['config']['parity_spot_check_all_ranks'], d['also']['cfg5_mixed']['ms_per_batch'], d['also']['cfg5_mixed']['parity_one_image_per_size_rank0'])" >> gpurun_out/c6_matrix.log 2>&1
done
B2="python bench.py --steps 3 --warmup 3 --no-cpu --no-e2e --also none"
$B2 > gpurun_out/c6_plain_mma.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:k_preprocess_fused -s 3 -c 1 -o gpurun_out/c6_mma $B2 > gpurun_out/c6_ncu_mma.log 2>&1
tail -5 gpurun_out/c6_pytest.log; grep -v "fused_accepts" gpurun_out/c6_matrix.log
