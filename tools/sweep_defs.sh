#!/bin/bash
# development sweep ON the GPU box: rebuild libb200rt.so with each set of -D knobs and time the probe
# usage: tools/sweep_defs.sh "<defs 1>" "<defs 2>" ...
for defs in "$@"; do
  B200RT_NVCC_DEFS="$defs" python -c "from raytracing_course_hw_b200 import build; build.build_cuda(force=True)" || exit 1
  echo "defs [$defs]: $(python tools/profile_step.py --steps 3 --counters 2>&1 | tail -2 | head -1 | sed 's/.*node_tests/node_tests/' | cut -c1-330)"
  echo "     $(python tools/profile_step.py --steps 3 2>&1 | grep 'step 2')"
done
python -c "from raytracing_course_hw_b200 import build; build.build_cuda(force=True)"
