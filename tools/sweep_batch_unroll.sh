for u in 1 2; do
  B200RT_NVCC_DEFS="-DB200RT_EXT_UNROLL=$u" python -c "from raytracing_course_hw_b200 import build; build.build_cuda(force=True)" || exit 1
  for b in 4 6 8 12 16 24; do for l in 8 12; do
    echo "unroll $u batch $b leaf_t $l: $(B200RT_WF_BATCH=$b B200RT_WF_LEAF_T=$l python tools/profile_step.py --steps 3 --serial 2>&1 | grep 'step 2' | sed 's/.*| //')"
  done; done
done
python -c "from raytracing_course_hw_b200 import build; build.build_cuda(force=True)"
