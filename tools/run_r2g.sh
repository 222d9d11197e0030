# Round-2 run G (2 GPUs): NCCL correctness tests, then the bench at N = 2 (weak line + strong record + PRB at 2 GPUs)
set -x
nvidia-smi -L
timeout 600 python -m pytest tests/test_gpu_multi.py tests/test_gpu_parity.py -m gpu -q -k "two_rank or alias or prb_sample_shards" > gpurun_out/r2g_pytest.log 2>&1; tail -5 gpurun_out/r2g_pytest.log
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 5 --warmup 3 > gpurun_out/r2g_bench2.log 2>&1; tail -1 gpurun_out/r2g_bench2.log | cut -c1-400
