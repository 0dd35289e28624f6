#!/bin/bash
# one build->measure iteration on the GPU box: parity tests, short bench, optional ncu full capture of the main env_step launch
TAG=${1:-iter}
timeout 1200 python -m pytest tests -m gpu -q --tb=short -x 2>&1 | tail -25 | cut -c1-300 > gpurun_out/pytest_$TAG.log
tail -5 gpurun_out/pytest_$TAG.log
timeout 600 python bench.py --steps 20 --warmup 3 --no-cpu-baseline 2>gpurun_out/bench_$TAG.err | tail -1 > gpurun_out/bench_$TAG.log
python -c "
import json; d=json.load(open('gpurun_out/bench_$TAG.log')); print('value',d['value'],'ms/step', d['ms_per_step'],'kernel_ms', d['roofline']['kernel_ms'],'frac', d['roofline']['frac'],'e2e', d['e2e']['value'])"
if [ "$2" != "noncu" ]; then
timeout 900 ncu --set full --clock-control none --import-source on -k regex:env_step_kernel -s 5 -c 1 -f -o gpurun_out/prof_env_step_$TAG python bench.py --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_$TAG.log 2>&1
tail -1 gpurun_out/ncu_$TAG.log | cut -c1-100
fi
