ncu --set full --import-source on --clock-control none --kernel-name regex:k_wf_trace --launch-skip 2 --launch-count 1 -o gpurun_out/r01_v4_trace_bounce2 -f python tools/profile_render.py bedroom 1 16 > gpurun_out/ncu4.log 2>&1
ls -la gpurun_out/*.ncu-rep
