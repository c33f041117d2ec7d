#!/bin/bash
# development sweep ON the GPU box over run-time knobs at the BENCH size: tools/sweep_env_bench.sh "A=1 B=2" "A=2" ...
for kv in "$@"; do
  echo "[$kv] $(env $kv python bench.py --steps 4 --warmup 3 --no-cpu-baseline 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('value %.1f e2e %.1f ms/step %.1f launches %d' % (d['value'], d['e2e']['value'], d['ms_per_step'], d['gpu_launches']))")"
done
