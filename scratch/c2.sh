#!/bin/bash
# GPU call 2: occupancy/partition matrix for the fused kernel + ncu full captures
mkdir -p gpurun_out
B="python bench.py --steps 30 --warmup 5 --no-cpu --no-e2e --also none"
for mma in 1 0; do for per in "" 1 2 3 4 7; do
  echo "== MMA=$mma CTAS_PER_SM=$per" >> gpurun_out/c2_matrix.log
  TCB_DEBUG=1 TCB_FUSED_MMA=$mma TCB_FUSED_CTAS_PER_SM=$per timeout 120 $B 2>> gpurun_out/c2_matrix.log | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(d['ms_per_step'], d['roofline']['frac'], d['config']['parity_spot_check_all_ranks'])" >> gpurun_out/c2_matrix.log 2>&1
done; done
B2="python bench.py --steps 3 --warmup 3 --no-cpu --no-e2e --also none"
$B2 > gpurun_out/c2_plain_mma.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:k_preprocess_fused -s 3 -c 2 -o gpurun_out/c2_mma $B2 > gpurun_out/c2_ncu_mma.log 2>&1
TCB_FUSED_MMA=0 $B2 > gpurun_out/c2_plain_dp4a.log 2>&1 && TCB_FUSED_MMA=0 ncu --set full --clock-control none --import-source on -k regex:k_preprocess_fused -s 3 -c 2 -o gpurun_out/c2_dp4a $B2 > gpurun_out/c2_ncu_dp4a.log 2>&1
cat gpurun_out/c2_matrix.log
