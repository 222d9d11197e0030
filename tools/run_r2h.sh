set -x
python tools/profile_render.py bedroom 2 64 > gpurun_out/r2h_plain.log 2>&1 && \
ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,smsp__thread_inst_executed_per_inst_executed.ratio,smsp__issue_active.avg.pct_of_peak_sustained_active --clock-control none -k regex:k_wf -s 58 -c 58 --csv --log-file gpurun_out/r2h_launches.csv python tools/profile_render.py bedroom 2 64 > gpurun_out/r2h_ncu_list.log 2>&1
