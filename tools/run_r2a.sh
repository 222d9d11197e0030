# Round-2 run A: full GPU test suite (incl. the config-scale parity tests), bench with the new records, PRB ncu evidence
set -x
nvidia-smi -L
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/r2a_pytest.log 2>&1; tail -5 gpurun_out/r2a_pytest.log
python bench.py --steps 5 --warmup 3 > gpurun_out/r2a_bench.log 2>&1; tail -1 gpurun_out/r2a_bench.log | cut -c1-1500
python tools/profile_prb.py 2 > gpurun_out/r2a_prb_plain.log 2>&1 && \
ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,lts__t_sectors_op_red.sum,lts__t_sectors_op_atom.sum --clock-control none -c 400 --csv --log-file gpurun_out/r2a_prb_launches.csv \
    python tools/profile_prb.py 2 > gpurun_out/r2a_prb_ncu_list.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:k_wf_shade -s 25 -c 1 -o gpurun_out/r2a_prb_adj_shade python tools/profile_prb.py 2 > gpurun_out/r2a_prb_ncu_full.log 2>&1
ls -la gpurun_out | tail -12
