#!/bin/bash
mkdir -p gpurun_out
rm -f gpurun_out/c18_*
B="python bench.py --steps 50 --warmup 5 --no-cpu --no-e2e"
run() { echo "== $1" >> gpurun_out/c18_matrix.log; env $1 TCB_DEBUG=1 timeout 120 $B --also none 2> gpurun_out/c18_err.log | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(d['ms_per_step'], d['roofline']['frac'], d['config']['parity_spot_check_all_ranks'])" >> gpurun_out/c18_matrix.log 2>&1; grep "fused class" gpurun_out/c18_err.log | sort | uniq -c | head -8 >> gpurun_out/c18_matrix.log; }
run "TCB_FUSED_WS=1"
run "TCB_FUSED_WS=2"
export TCB_FUSED_WS=1
B2="python bench.py --steps 2 --warmup 3 --no-cpu --no-e2e --also none"
timeout 200 $B2 > gpurun_out/c18_plain2.log 2>&1 && timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_preprocess_fused_ws -s 3 -c 1 -o gpurun_out/c18_ws $B2 > gpurun_out/c18_ncu.log 2>&1
cat gpurun_out/c18_matrix.log; tail -3 gpurun_out/c18_ncu.log
