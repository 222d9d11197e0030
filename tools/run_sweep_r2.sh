# sweep of the trace kernel's vote / refill thresholds on the bedroom workload
for r in 4 8 12 16 24; do echo "refill=$r tri=10 $(B2_REFILL_THRESH=$r B2_TRI_THRESH=10 python tools/profile_render.py bedroom 3 64 2>&1 | grep 'launch 2')"; done
for t in 6 14 18 24; do echo "refill=12 tri=$t $(B2_REFILL_THRESH=12 B2_TRI_THRESH=$t python tools/profile_render.py bedroom 3 64 2>&1 | grep 'launch 2')"; done
