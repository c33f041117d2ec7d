B200RT_NVCC_DEFS="-DB200RT_DBG_SORT" python -c "from raytracing_course_hw_b200 import build; build.build_cuda(force=True)" || exit 1
echo "none: $(python tools/profile_step.py --steps 3 --serial 2>&1 | grep 'step 2')"
for mode in 0 1 2 3; do for bits in 3 5 7; do
  echo "mode $mode bits $bits: $(B200RT_DBG_SORT_MODE=$mode B200RT_DBG_SORT_BITS=$bits python tools/profile_step.py --steps 3 --serial 2>&1 | grep 'step 2')"
done; done
python -c "from raytracing_course_hw_b200 import build; build.build_cuda(force=True)"
