#!/bin/bash
N=$1
mkdir -p gpurun_out
t0=$(date +%s)
timeout 1500 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus $N > gpurun_out/final_n$N.json 2> gpurun_out/final_n$N.err
echo "rc=$? secs=$(( $(date +%s) - t0 ))" > gpurun_out/final_n$N.rc
if [ "$N" = "2" ]; then timeout 600 python -m pytest tests/test_multi_gpu.py -m gpu -q > gpurun_out/final_n2_pytest.log 2>&1; tail -2 gpurun_out/final_n2_pytest.log; fi
cat gpurun_out/final_n$N.rc; tail -3 gpurun_out/final_n$N.err; head -c 300 gpurun_out/final_n$N.json
