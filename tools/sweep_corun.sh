# development: can the latency-bound shade of one wave run beside the issue-bound persistent traversal of another,
# if the traversal kernels leave part of each SM's register file free?  (B200RT_DEV build: env knobs are live)
B200RT_NVCC_DEFS="-DB200RT_DEV" python -c "from raytracing_course_hw_b200 import build; build.build_cuda(force=True)" || exit 1
for pipe in 2 3; do for cap in 0 8 7 6 5 4; do
  echo "pipe=$pipe blocks_per_sm_cap=$cap: $(B200RT_WF_PIPE=$pipe B200RT_WF_BLOCKS_PER_SM=$cap python tools/profile_step.py --width 3840 --height 2160 --spp 8 --steps 3 | grep 'step 2' | cut -c1-120)"
done; done
python -c "from raytracing_course_hw_b200 import build; build.build_cuda(force=True)"
