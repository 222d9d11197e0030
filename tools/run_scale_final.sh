N=${1:-2}
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29521 bench.py --gpus $N --steps 5 --warmup 3 > gpurun_out/final_bench_${N}gpu.log 2>&1; tail -1 gpurun_out/final_bench_${N}gpu.log | cut -c1-300
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29523 bench.py --impl reference --gpus $N --steps 1 --warmup 1 > gpurun_out/final_bench_ref_${N}gpu.log 2>&1; tail -1 gpurun_out/final_bench_ref_${N}gpu.log | cut -c1-200
