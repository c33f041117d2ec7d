for b in 4 8 16 32; do for l in 4 8 12 16 24; do echo "batch $b leaf $l: $(B200RT_WF_BATCH=$b B200RT_WF_LEAF_T=$l B200RT_WF_LEAF_T_LIGHT=$l python tools/profile_step.py --steps 3 | grep 'step 2')"; done; done
for c in 2 4 6; do echo "blocks/SM $c: $(B200RT_WF_BLOCKS_PER_SM=$c python tools/profile_step.py --steps 3 | grep 'step 2')"; done
