for k in 0 2 3 4 6 8; do echo "light_top $k: $(B200RT_LIGHT_TOP=$k python tools/profile_step.py --steps 3 --counters | tail -2 | head -1 | sed 's/.*light_queries/lq/')"; done
for k in 0 4 8; do echo "light_top $k: $(B200RT_LIGHT_TOP=$k python tools/profile_step.py --steps 3 | grep 'step 2')"; done
python -m pytest tests -m gpu -x -q 2>&1 | tail -3
