python -m pytest tests -m gpu -q -k "prb or forward or front_door or optimisation or c3" 2>&1 | tail -4
python tools/profile_prb.py 3 2>&1 | tail -2
B2_HIT_HISTORY_GB=0 python tools/profile_prb.py 3 2>&1 | tail -1
