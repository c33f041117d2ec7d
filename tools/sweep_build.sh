# development sweep: rebuild the CUDA library with different compile-time knobs ON the GPU box
for defs in "" "-DB200RT_EXT_MINB=8" "-DB200RT_EXT_MINB=9"; do
  B200RT_NVCC_DEFS="$defs" python -c "from raytracing_course_hw_b200 import build; build.build_cuda(force=True)" || exit 1
  echo "defs [$defs]: $(python tools/profile_step.py --steps 3 --counters | tail -2 | head -1 | cut -c1-420)"
  echo "     $(python tools/profile_step.py --steps 3 | grep 'step 2')"
done
python -c "from raytracing_course_hw_b200 import build; build.build_cuda(force=True)"
