for lib in "$@"; do
  PARC_B200_LIB=/root/repo/$lib python bench_optimizer.py --cpu-motions 0 2>/dev/null | tail -1 | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('$lib', d['ms_per_eval'])"
done
