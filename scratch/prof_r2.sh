#!/bin/bash
# round-2 profiles: launch list of one bench pass, then --set full captures of the hot kernels
mkdir -p gpurun_out
B="python bench.py --steps 2 --warmup 3 --no-cpu --no-e2e --also cfg2_serialize,cfg4_deserialize,cfg5_mixed,bf16"
$B > gpurun_out/r2f_plain.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/r2f_launches.csv $B > gpurun_out/r2f_ncu_launches.log 2>&1
B2="python bench.py --steps 2 --warmup 3 --no-cpu --no-e2e --also none"
$B2 > gpurun_out/r2f_plain2.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:k_preprocess_fused_mma -s 3 -c 1 -o gpurun_out/r2f_fused_mma $B2 > gpurun_out/r2f_ncu_mma.log 2>&1
B3="python bench.py --steps 2 --warmup 3 --no-cpu --no-e2e --also cfg4_deserialize,cfg2_serialize,bf16"
$B3 > gpurun_out/r2f_plain3.log 2>&1 && ncu --set full --clock-control none -k regex:k_des_gather\|k_des_chase\|k_des_vote\|k_bytes_serialize\|k_bf16 -c 24 -o gpurun_out/r2f_bytes $B3 > gpurun_out/r2f_ncu_bytes.log 2>&1
ls -la gpurun_out/
