#!/bin/bash
# round-end evidence on one B200: parity tests, smoke, both bench arms, ncu launch list of the bench command
TAG=${1:-r1}
python -m pytest tests -m gpu -q --tb=short 2>&1 | tail -5 | cut -c1-300
python __graft_entry__.py smoke 2>&1 | tail -2
python bench.py --impl reference --steps 10 --warmup 3 2>/dev/null | tail -1 > gpurun_out/bench_ref_$TAG.json
python bench.py 2>gpurun_out/bench_$TAG.err | tail -1 > gpurun_out/bench_$TAG.json
cut -c1-600 gpurun_out/bench_$TAG.json
ncu --metrics gpu__time_duration.sum --clock-control none -k regex:"env_step|env_reset|fail_rate|zero_int|sim_standin" -c 80 --csv \
    --log-file gpurun_out/launches_$TAG.csv python bench.py --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_l.log 2>&1
tail -1 gpurun_out/ncu_l.log | cut -c1-120
