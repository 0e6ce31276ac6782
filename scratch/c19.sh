#!/bin/bash
mkdir -p gpurun_out
rm -f gpurun_out/c19_*
(timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/c19_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/c19_pytest.log)
timeout 300 python scratch/cfg5_time.py > gpurun_out/c19_cfg5.log 2>&1
timeout 300 python bench.py --steps 20 --warmup 3 --no-cpu --no-e2e --also cfg5_mixed > gpurun_out/c19_bench.json 2> gpurun_out/c19_bench.err
tail -4 gpurun_out/c19_pytest.log; cat gpurun_out/c19_cfg5.log | tail -11; python -c "
import json; d=json.load(open('gpurun_out/c19_bench.json')); print(d['ms_per_step'], d['roofline']['frac'], d['also']['cfg5_mixed']['ms_per_batch'], d['also']['cfg5_mixed']['roofline']['frac'], d['also']['cfg5_mixed']['parity_one_image_per_size_rank0'])"
