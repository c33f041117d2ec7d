#!/bin/bash
# development sweep ON the GPU box: rebuild libb200rt.so with each set of -D knobs; time the 1080p probe with one
# wave in flight (exclusive stage times) and with the default overlap
# usage: tools/sweep_defs_serial.sh "<defs 1>" "<defs 2>" ...
for defs in "$@"; do
  B200RT_NVCC_DEFS="$defs" python -c "from raytracing_course_hw_b200 import build; build.build_cuda(force=True)" || exit 1
  echo "defs [$defs]"
  echo "   serial : $(python tools/profile_step.py --steps 3 --serial 2>&1 | grep 'step 2')"
  echo "   overlap: $(python tools/profile_step.py --steps 3 2>&1 | grep 'step 2')"
done
python -c "from raytracing_course_hw_b200 import build; build.build_cuda(force=True)"
