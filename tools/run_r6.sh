python -m pytest tests -m gpu -x -q 2>&1 | tail -5 > gpurun_out/pytest_gpu6.log; tail -2 gpurun_out/pytest_gpu6.log
(
for b in 0 1; do echo "== bedroom builder=$b"; B2_BVH_BUILDER=$b python tools/profile_render.py bedroom 2 16; done
for b in 0 1; do echo "== tris builder=$b"; B2_BVH_BUILDER=$b python tools/profile_render.py tris 2 4; done
) > gpurun_out/sweep6.log 2>&1
grep -E "==|launch 1|build" gpurun_out/sweep6.log | cut -c1-400
python bench.py --steps 5 --warmup 3 > gpurun_out/bench6.log 2>&1; tail -1 gpurun_out/bench6.log
