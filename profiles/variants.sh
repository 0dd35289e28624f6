#!/bin/bash
# A/B runs of library variants built into scratch/ (same sources, other -D tuning constants): bench kernel time per variant
for lib in "$@"; do
  PARC_B200_LIB=/root/repo/$lib timeout 600 python bench.py --steps 20 --warmup 3 --no-cpu-baseline 2>/dev/null | tail -1 > gpurun_out/variant.json
  python -c "
import json; d=json.load(open('gpurun_out/variant.json')); print('$lib', 'kernel_ms %.4f' % d['roofline']['kernel_ms'], 'step_ms %.4f' % d['ms_per_step'], 'frac %.3f' % d['roofline']['frac'], 'e2e_ms %.3f' % d['e2e']['ms_per_step'])"
done
