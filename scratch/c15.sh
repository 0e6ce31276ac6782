#!/bin/bash
mkdir -p gpurun_out
(timeout 600 python -m pytest tests/test_preprocess_gpu.py tests/test_multi_gpu.py -m gpu -x -q > gpurun_out/c15_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/c15_pytest.log)
B="python bench.py --steps 30 --warmup 5 --no-cpu --no-e2e"
run() { echo "== $1" >> gpurun_out/c15_matrix.log; env $1 TCB_DEBUG=1 timeout 120 $B --also cfg5_mixed 2>> gpurun_out/c15_matrix.log | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(d['ms_per_step'], d['roofline']['frac'], d['config']['parity_spot_check_all_ranks'], d['also']['cfg5_mixed']['ms_per_batch'], d['also']['cfg5_mixed']['parity_one_image_per_size_rank0'])" >> gpurun_out/c15_matrix.log 2>&1; }
run "X=1"
run "TCB_MMA_W8=1"
run "TCB_MMA_SPLIT=0"
run "TCB_FUSED_CTAS_PER_SM=3"
run "TCB_FUSED_CTAS_PER_SM=5 TCB_MMA_W8=1"
B2="python bench.py --steps 3 --warmup 3 --no-cpu --no-e2e --also none"
$B2 > gpurun_out/c15_plain.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:k_preprocess_fused -s 3 -c 1 -o gpurun_out/c15_mma $B2 > gpurun_out/c15_ncu.log 2>&1
tail -4 gpurun_out/c15_pytest.log; grep -v fused_accepts gpurun_out/c15_matrix.log
