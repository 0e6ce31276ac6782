#!/bin/bash
mkdir -p gpurun_out
rm -f gpurun_out/c17_*
(TCB_FUSED_WS=1 timeout 600 python -m pytest tests/test_preprocess_gpu.py tests/test_multi_gpu.py -m gpu -x -q > gpurun_out/c17_pytest_ws.log 2>&1; echo "pytest rc=$?" >> gpurun_out/c17_pytest_ws.log)
B="python bench.py --steps 50 --warmup 5 --no-cpu --no-e2e"
run() { echo "== $1" >> gpurun_out/c17_matrix.log; env $1 TCB_DEBUG=1 timeout 120 $B --also cfg5_mixed 2> gpurun_out/c17_err.log | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(d['ms_per_step'], d['roofline']['frac'], d['config']['parity_spot_check_all_ranks'], d['also']['cfg5_mixed']['ms_per_batch'], d['also']['cfg5_mixed']['parity_one_image_per_size_rank0'])" >> gpurun_out/c17_matrix.log 2>&1; grep "fused class" gpurun_out/c17_err.log | sort | uniq -c | head -8 >> gpurun_out/c17_matrix.log; }
run "TCB_FUSED_WS=0"
run "TCB_FUSED_WS=1"
run "TCB_FUSED_WS=2"
tail -5 gpurun_out/c17_pytest_ws.log; cat gpurun_out/c17_matrix.log
