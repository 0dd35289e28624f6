for args in "--fail-rate-sync 0 --steps 30" "--fail-rate-sync 10 --steps 30" "--fail-rate-sync 10 --steps 200" "--fail-rate-sync 1 --steps 100"; do
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --warmup 5 --no-extra --no-cpu-baseline $args 2>/dev/null | tail -1 | python -c "
import json,sys; d=json.loads(sys.stdin.read()); print('$args', d['value'], d['ms_per_step'], d['host_enqueue_ms_per_step'], d['e2e']['ms_per_step'])"
done
