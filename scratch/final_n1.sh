#!/bin/bash
mkdir -p gpurun_out
rm -f gpurun_out/final_*
t0=$(date +%s)
timeout 1500 python bench.py > gpurun_out/final_n1.json 2> gpurun_out/final_n1.err; echo "rc=$? secs=$(( $(date +%s) - t0 ))" > gpurun_out/final_n1.rc
t0=$(date +%s)
timeout 900 python bench.py --impl reference > gpurun_out/final_ref_n1.json 2> gpurun_out/final_ref_n1.err; echo "rc=$? secs=$(( $(date +%s) - t0 ))" >> gpurun_out/final_n1.rc
cat gpurun_out/final_n1.rc; tail -3 gpurun_out/final_n1.err; head -c 600 gpurun_out/final_ref_n1.json
