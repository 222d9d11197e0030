cd mitsuba3_experiments_b200/csrc; cp libb2path.so /tmp/lib_orig.so
for mb in 3 5; do cp alt/libb2path_mb$mb.so libb2path.so; echo "== trace minblocks=$mb"; (cd ../..; python tools/profile_render.py bedroom 3 64 2>&1 | grep "launch 2"; python tools/profile_render.py tris 2 8 2>&1 | grep "launch 1"); done
cp /tmp/lib_orig.so libb2path.so; echo "== default (4)"; (cd ../..; python tools/profile_render.py bedroom 3 64 2>&1 | grep "launch 2")
