#!/bin/bash
mkdir -p gpurun_out
B="python bench.py --steps 2 --warmup 3 --no-cpu --no-e2e --also cfg5_mixed"
$B > gpurun_out/r2f_cfg5_plain.log 2>&1 && ncu --set full --clock-control none --import-source on -k 'regex:^k_preprocess_fused' -s 6 -c 10 -o gpurun_out/r2f_cfg5 $B > gpurun_out/r2f_cfg5_ncu.log 2>&1
tail -3 gpurun_out/r2f_cfg5_ncu.log
