python -m pytest tests -m gpu -x -q 2>&1 | tail -15 > gpurun_out/pytest_gpu2.log; tail -3 gpurun_out/pytest_gpu2.log
(
echo "== greedy"; B2_BVH_COLLAPSE=greedy python tools/profile_render.py bedroom 3 64
for cp in 0.2 0.3 0.45 0.7; do echo "== dp c_prim=$cp"; B2_BVH_CPRIM=$cp python tools/profile_render.py bedroom 3 64; done
echo "== tris greedy"; B2_BVH_COLLAPSE=greedy python tools/profile_render.py tris 2 8
echo "== tris dp"; python tools/profile_render.py tris 2 8
echo "== e2e"; python tools/diag_e2e.py
) > gpurun_out/sweep2.log 2>&1
grep -E "==|launch 2|launch 1|wall" gpurun_out/sweep2.log
