#!/bin/bash
mkdir -p gpurun_out
rm -f gpurun_out/c20_*
(timeout 900 python -m pytest tests/test_postprocess.py tests/test_preprocess_gpu.py -m gpu -x -q > gpurun_out/c20_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/c20_pytest.log)
timeout 300 python bench.py --steps 20 --warmup 3 --no-cpu --no-e2e --also cfg5_mixed > gpurun_out/c20_bench.json 2> gpurun_out/c20_bench.err
tail -12 gpurun_out/c20_pytest.log; python -c "
import json; d=json.load(open('gpurun_out/c20_bench.json')); print(d['ms_per_step'], d['roofline']['frac'], d['also']['cfg5_mixed']['ms_per_batch'], d['also']['cfg5_mixed']['roofline']['frac'], d['also']['cfg5_mixed']['parity_one_image_per_size_rank0'])"
