# Round-2 run O: C5 (bedroom 3840x2160, 1024 spp, 2 passes x 512) on ONE GPU: the partner of the 8-GPU number
set -x
python bench.py --gpus 1 --steps 2 --warmup 3 --workload bedroom4k --no-prb --no-cpu-baseline > gpurun_out/r2o_bench4k_1.log 2>&1; tail -1 gpurun_out/r2o_bench4k_1.log | cut -c1-300
