#!/bin/bash
# development sweep ON the GPU box over run-time knobs: tools/sweep_env.sh "A=1 B=2" "A=2" ...
for kv in "$@"; do
  echo "[$kv] $(env $kv python tools/profile_step.py --spp ${SPP:-8} --steps 3 --counters 2>&1 | tail -2 | head -1 | sed 's/.*extend_ms/extend_ms/')  $(env $kv python tools/profile_step.py --spp ${SPP:-8} --steps 3 2>&1 | grep 'step 2')"
done
