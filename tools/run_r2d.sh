# Round-2 run D: full GPU suite on the rewritten wavefront + ncu of the trace kernel
set -x
timeout 1500 python -m pytest tests -m gpu -q --durations=8 > gpurun_out/r2d_pytest.log 2>&1; tail -15 gpurun_out/r2d_pytest.log
python tools/profile_render.py bedroom 2 64 > gpurun_out/r2d_plain.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:k_wf_trace -s 20 -c 1 -o gpurun_out/r2d_trace python tools/profile_render.py bedroom 2 64 > gpurun_out/r2d_ncu_full.log 2>&1
tail -3 gpurun_out/r2d_plain.log
