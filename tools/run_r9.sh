(
for t in 4 6 8 12 16; do echo "== tri_thresh=$t"; B2_BVH_BUILDER=1 B2_TRI_THRESH=$t python tools/profile_render.py bedroom 2 64 | grep "launch 1"; done
for s in 3 4 5 6 8 12; do echo "== slice=$s"; B2_BVH_BUILDER=1 B2_TRACE_SLICE=$s python tools/profile_render.py bedroom 2 64 | grep "launch 1"; done
for c in 0.6 1.0 1.5; do echo "== c_node=$c"; B2_BVH_BUILDER=1 B2_BVH_CNODE=$c python tools/profile_render.py bedroom 2 64 | grep "launch 1"; done
) > gpurun_out/sweep9.log 2>&1
cat gpurun_out/sweep9.log
