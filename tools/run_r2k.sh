# Round-2 run K (8 GPUs): bench at N = 8 (weak line + strong record + PRB at 8 GPUs), C5 (bedroom 4K, 1024 spp, strong) at N = 8
set -x
nvidia-smi -L | wc -l
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29521 bench.py --gpus 8 --steps 5 --warmup 3 > gpurun_out/r2k_bench8.log 2>&1; tail -1 gpurun_out/r2k_bench8.log | cut -c1-300
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29522 bench.py --gpus 8 --steps 2 --warmup 3 --workload bedroom4k --no-prb > gpurun_out/r2k_bench4k_8.log 2>&1; tail -1 gpurun_out/r2k_bench4k_8.log | cut -c1-300
