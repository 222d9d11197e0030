(
for s in 7 8 10; do for t in 10 12; do echo "== slice=$s tri=$t"; B2_BVH_BUILDER=1 B2_TRACE_SLICE=$s B2_TRI_THRESH=$t python tools/profile_render.py bedroom 2 64 | grep "launch 1"; done; done
for s in 5 8; do echo "== tris slice=$s tri=12"; B2_TRACE_SLICE=$s B2_TRI_THRESH=12 python tools/profile_render.py tris 2 8 | grep "launch 1"; done
for s in 5 8; do echo "== cornell slice=$s tri=12"; B2_TRACE_SLICE=$s B2_TRI_THRESH=12 python tools/profile_render.py cornell 3 16 | grep "launch 2"; done
) > gpurun_out/sweep10.log 2>&1
cat gpurun_out/sweep10.log | paste - - | cut -c1-100
