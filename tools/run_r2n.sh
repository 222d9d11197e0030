# Round-2 run N: the other workloads once, for the record (C1, C4, scene.xml palette, alias-table mode)
set -x
python bench.py --steps 5 --warmup 3 --workload cornell --no-prb > gpurun_out/r2n_bench_cornell.log 2>&1; tail -1 gpurun_out/r2n_bench_cornell.log | cut -c1-200
python bench.py --steps 5 --warmup 3 --workload bedroom_xml --no-prb > gpurun_out/r2n_bench_xml.log 2>&1; tail -1 gpurun_out/r2n_bench_xml.log | cut -c1-200
python bench.py --steps 5 --warmup 3 --env-warp alias --no-prb --no-cpu-baseline > gpurun_out/r2n_bench_alias.log 2>&1; tail -1 gpurun_out/r2n_bench_alias.log | cut -c1-200
python bench.py --steps 3 --warmup 3 --workload soup --no-prb --no-cpu-baseline > gpurun_out/r2n_bench_soup.log 2>&1; tail -1 gpurun_out/r2n_bench_soup.log | cut -c1-200
