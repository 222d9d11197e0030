python -m pytest tests -m gpu -x -q 2>&1 | tail -15 > gpurun_out/pytest_gpu5.log; tail -3 gpurun_out/pytest_gpu5.log
(
for b in 0 1; do for pf in 0 1; do echo "== builder=$b prefetch=$pf"; B2_BVH_BUILDER=$b B2_PREFETCH=$pf python tools/profile_render.py bedroom 3 64; done; done
for r in 4 16 25; do echo "== ploc radius=$r"; B2_PLOC_RADIUS=$r python tools/profile_render.py bedroom 3 64; done
for cp in 0.2 0.5; do echo "== ploc c_prim=$cp"; B2_BVH_CPRIM=$cp python tools/profile_render.py bedroom 3 64; done
echo "== tris lbvh"; B2_BVH_BUILDER=0 python tools/profile_render.py tris 2 8
echo "== tris ploc"; python tools/profile_render.py tris 2 8
echo "== cornell ploc"; python tools/profile_render.py cornell 3 16
) > gpurun_out/sweep5.log 2>&1
grep -E "==|launch 2|launch 1|build|Error|error" gpurun_out/sweep5.log | cut -c1-330
