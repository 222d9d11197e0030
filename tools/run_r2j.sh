# Round-2 run J: new PRB gradient targets + PRB ncu evidence on the v6 kernels
set -x
timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -q -k "rough_bsdf_gradient or prb_ or front_door or optimisation" > gpurun_out/r2j_pytest.log 2>&1; tail -12 gpurun_out/r2j_pytest.log
python tools/profile_prb.py 2 > gpurun_out/r2j_prb_plain.log 2>&1 && \
ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,lts__t_sectors_op_red.sum,smsp__thread_inst_executed_per_inst_executed.ratio,smsp__issue_active.avg.pct_of_peak_sustained_active --clock-control none -k regex:k_wf -s 82 -c 60 --csv --log-file gpurun_out/r2j_prb_launches.csv \
    python tools/profile_prb.py 2 > gpurun_out/r2j_prb_ncu_list.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:k_wf_shade -s 25 -c 1 -o gpurun_out/r2j_prb_adj_shade python tools/profile_prb.py 2 > gpurun_out/r2j_prb_ncu_full.log 2>&1
tail -2 gpurun_out/r2j_prb_plain.log
