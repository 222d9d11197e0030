(
for m in 1 2 4; do echo "== shade mult=$m"; B2_BVH_BUILDER=1 B2_SHADE_GRID_MULT=$m python tools/profile_render.py bedroom 3 64; done
echo "== cornell"; python tools/profile_render.py cornell 3 16
) > gpurun_out/sweep8.log 2>&1
grep -E "==|launch 2" gpurun_out/sweep8.log | cut -c1-300
