python -m pytest tests -m gpu -x -q 2>&1 | tail -8 > gpurun_out/pytest_gpu17.log; tail -3 gpurun_out/pytest_gpu17.log
python __graft_entry__.py smoke 2>&1 | tail -2
(
echo "== C4 tris 1920x1080 128spp"; python tools/profile_render.py tris 2 128 | grep -E "build|launch 1"
echo "== cornell C1"; python tools/profile_render.py cornell 3 16 | grep -E "launch 2"
) > gpurun_out/sweep17.log 2>&1; cat gpurun_out/sweep17.log | cut -c1-400
