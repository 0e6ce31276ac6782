#!/bin/bash
timeout 600 python bench.py --no-cpu --also bf16,cfg2_serialize,cfg1_latency 2>/dev/null | python -c "
import sys,json; d=json.loads(sys.stdin.read()); a=d['also']
print(d['ms_per_step'], d['roofline']['frac'], d['clocks'])
print('e2e', d['e2e']['value'], d['e2e']['frac_of_h2d_ceiling'], d['e2e']['from_pageable']['value'])
print('cfg2', a['cfg2_serialize']['us_per_launch'], a['cfg2_serialize']['roofline']['frac'])
print('bf16', a['bf16']['serialize_us'], a['bf16']['deserialize_us'], a['bf16']['roofline']['frac'], a['bf16']['roofline_deserialize']['frac'])
print('cfg1', a['cfg1_latency']['value'])"
