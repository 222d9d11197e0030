python -m pytest tests -m gpu -x -q 2>&1 | tail -20 > gpurun_out/pytest_gpu20.log; tail -20 gpurun_out/pytest_gpu20.log
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29541 bench.py --gpus 2 --steps 5 --warmup 3 > gpurun_out/bench_2gpu_s.log 2>&1; tail -1 gpurun_out/bench_2gpu_s.log | cut -c1-220
