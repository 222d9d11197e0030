set -x
python tools/profile_render.py bedroom 2 64 > gpurun_out/r2i_plain.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:k_wf_shade -s 45 -c 1 -o gpurun_out/r2i_shade_b1 python tools/profile_render.py bedroom 2 64 > gpurun_out/r2i_ncu2.log 2>&1
ls -la gpurun_out
