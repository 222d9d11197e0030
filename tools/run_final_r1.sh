# Round-1 final evidence: smoke, plain bench, ncu launch list of the same bench command, one --set full capture of k_wf_trace
set -x
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/final_smoke.log 2>&1; tail -2 gpurun_out/final_smoke.log
python bench.py > gpurun_out/final_bench.log 2>&1; tail -1 gpurun_out/final_bench.log | cut -c1-600
python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/final_bench_ref.log 2>&1; tail -1 gpurun_out/final_bench_ref.log | cut -c1-300
python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-prb > gpurun_out/final_bench_short.log 2>&1 && \
ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -c 2000 --csv --log-file gpurun_out/final_launches.csv \
    python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-prb > gpurun_out/final_ncu_list.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:k_wf_trace -s 40 -c 1 -o gpurun_out/final_trace python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-prb > gpurun_out/final_ncu_full.log 2>&1
ls -la gpurun_out | tail -8
